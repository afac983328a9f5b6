"""CPU oracle for the weather-MLP DDP training step -- TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU, the arithmetic of the reference's hot path
(`/root/reference/jobs/preprocess.py:23-41` and
`/root/reference/jobs/train_lightning_ddp.py:16-88,117-123`).  It is the checker
for the CUDA path, the generator of `tests/golden/`, and the timed CPU baseline
in `bench.py` (`cpu_baseline` leg and `--impl reference`).  Nothing in the
product package (`wxddp`) may import it; only `tests/`, `__graft_entry__.smoke()`
and `bench.py` do.

Parity status (see DESIGN.md "Oracle"):
  * train step (model, loss, backward, Adam, dataset dtypes): PINNED -- the
    golden vectors in `tests/golden/` were produced by executing the reference's
    own `WeatherClassifier` / `WeatherDataset` classes from
    `/root/reference/jobs/train_lightning_ddp.py` in the authoring container
    (with thin stand-ins for the absent `pytorch_lightning` / `mlflow` shells,
    `tests/golden/make_golden.py`), and `oracle/train_ref.py` is checked against
    them bit-for-bit in `tests/test_oracle.py`.
  * Lightning's loop (sampler injection, metric sync, checkpoint call order) and
    Spark's `mean`/`stddev` (`preprocess.py:33`): PARITY UNPINNED -- neither
    pytorch-lightning 2.1.0 nor pyspark 3.4.1 (+JVM) is installable here and the
    reference holds no tests or golden vectors; these parts restate the
    libraries' published behaviour and are anchored on the reference's call sites.
"""
