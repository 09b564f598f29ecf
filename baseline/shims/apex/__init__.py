"""Stand-in for NVIDIA apex, which is not installed in this image. Used ONLY by the
reference arm of bench.py; every substitution is a PyTorch library op (documented in
DESIGN.md / BASELINE.md as "reference with torch-fused Adam instead of apex FusedAdam")."""
