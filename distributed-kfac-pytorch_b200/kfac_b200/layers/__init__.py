"""Per-layer K-FAC engine: module helpers (factor kernels) and layer state machines."""
