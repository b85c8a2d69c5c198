"""k8s_b200 -- the TfJob sync-PS hot path on B200 GPU ranks."""
__version__ = "0.1.0"
