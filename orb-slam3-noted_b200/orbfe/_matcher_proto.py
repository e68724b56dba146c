"""Prototypes of the matcher entry points of include/orbfe.h (filled as the kernels land)."""
PROTO = {}
