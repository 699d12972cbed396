// Diagnostic only: LD_PRELOAD this into the STOCK reference encoder to see what the mere presence of a CUDA context (its
// helper threads, its mappings) costs the host code.  Built and used by profiles/diag/host_slowdown.py.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
__attribute__((constructor)) static void go(void) {
  if (getenv("DIAG_NO_INIT")) return;
  cudaFree(0);
  if (getenv("DIAG_PIN_MB")) {            // optionally also page-lock some memory and keep a pinned allocation around
    void *p = nullptr;
    cudaMallocHost(&p, (size_t)atoi(getenv("DIAG_PIN_MB")) << 20);
  }
}
