#!/usr/bin/env python3
"""Drop-in for the arithmetic of the reference's `jobs/preprocess.py` (:23-41) on the GPU: label encode, column
statistics with the zero guard, z-score, written as the same `data.parquet/` directory (`*_norm` float64 columns +
int32 `label_encoded`) the trainer reads.  CSV parsing and the parquet writer stay on the host (pandas / pyarrow);
Spark itself is out of scope.

    python3 jobs/preprocess.py     # env: WX_RAW_CSV (/opt/spark/data/raw/weather.csv), WX_PROCESSED_DIR (/opt/spark/data/processed)
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from wxddp.etl import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
