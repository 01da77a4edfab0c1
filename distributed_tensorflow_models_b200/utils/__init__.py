"""Log-line formats, JSONL metrics, Chrome traces (``logs``), host/device pipelining (``pipeline``), TensorBoard
event files (``tfevents``)."""
