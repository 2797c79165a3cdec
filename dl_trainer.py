"""Entry point kept under the reference's file name; the implementation is gaussiank_sgd_b200/dl_trainer.py."""
from gaussiank_sgd_b200.dl_trainer import *  # noqa: F401,F403
from gaussiank_sgd_b200.dl_trainer import main

if __name__ == '__main__':
    main()
