"""Host-side loaders for the data formats that feed the hot path (FP8 KV-cache scaling factors)."""
