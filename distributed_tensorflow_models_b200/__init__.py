"""distributed_tensorflow_models_b200 — a B200-native (sm_100a) parameter-server data-parallel CNN training framework
with the capabilities of chenc10/distributed_TensorFlow_models.  See README.md / DESIGN.md / COVERAGE.md."""

__version__ = "0.1.0"
