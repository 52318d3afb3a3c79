"""Configuration / CLI (ref:139-165), data (ref:87-95), log formats (ref:114-117, :130-132),
checkpoints and the per-op timeline; ``ref`` = /root/reference/simple_distributed.py."""
