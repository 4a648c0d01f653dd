"""regcm_b200 — B200-native RegCM dynamical-core step behind a C ABI.

Package contents are only what the hot path needs: `csrc/` (CUDA kernels and
the C-ABI library), `capi` (ctypes mirror of include/regcm_dyn.h) and `synth`
(synthetic analytic DOMAIN/ICBC generator)."""
from . import capi, synth  # noqa: F401
