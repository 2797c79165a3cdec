"""Module kept under the reference's file name; the implementation is gaussiank_sgd_b200/ptb_reader.py."""
from gaussiank_sgd_b200.ptb_reader import *  # noqa: F401,F403
import gaussiank_sgd_b200.ptb_reader as _impl

globals().update({k: v for k, v in vars(_impl).items() if not k.startswith('__')})
