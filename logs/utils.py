"""Log-parsing helpers under the reference's module path (logs/utils.py:1-84): re-exported from the package."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiank_sgd_b200.utils.plotting import (get_real_title, read_data_from_log,  # noqa: E402,F401
                                               seconds_between_datetimestring)
