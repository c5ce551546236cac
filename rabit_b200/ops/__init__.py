"""Device ops: `emulate` drives the real kernels with N virtual ranks on one GPU."""
