"""One entry point per model x consistency mode (the reference's script names) over ``common.py``; ``train`` is the
generic one (any registry network); ``mobilenet_v1_train`` is the model zoo's own trainer."""
