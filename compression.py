"""Module kept under the reference's file name; the implementation is gaussiank_sgd_b200/compression.py."""
from gaussiank_sgd_b200.compression import *  # noqa: F401,F403
import gaussiank_sgd_b200.compression as _impl

globals().update({k: v for k, v in vars(_impl).items() if not k.startswith('__')})
