#!/usr/bin/env python3
"""Drop-in for the reference's `jobs/train_lightning_ddp.py`: same command line (none), same environment
contract, same files in and out -- the training step runs on B200 through libwxddp.so.

    python3 jobs/train_lightning_ddp.py           # one process per rank; exit code 0 on success

See wxddp/trainer.py (`main`) for the behaviour kept from the reference and the optional WX_* overrides.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from wxddp.module import WeatherClassifier, WeatherDataset  # noqa: E402,F401  (same public names as the reference)
from wxddp.trainer import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
