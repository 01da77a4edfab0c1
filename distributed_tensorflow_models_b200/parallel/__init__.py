"""Parameter-server data parallelism: flat shard layout, transports (thread cluster / gloo / NCCL baseline /
symmetric-memory NVLink kernels), the BSP / ASP / SSP engine, the SSP clock service and dedicated PS processes."""
