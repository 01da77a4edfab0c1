"""Checkpoints keyed by the reference's TensorFlow variable names: ``store`` (.npz container, state file, keep-N),
``tf_bundle`` (TensorFlow V2 bundles without TensorFlow), ``inspect`` (list / convert)."""
