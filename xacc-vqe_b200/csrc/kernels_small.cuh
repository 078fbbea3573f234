// Shared-memory-resident kernels for small registers (filled in by the batched sweep path).
#pragma once
#include "kernels.cuh"
