"""llm-mcp_b200 — B200-native Matryoshka truncate(+renormalise) for LLM-MCP's
``POST /v1/embeddings`` cloud branch (reference: core/internal/api/handlers.go:1891-1893,
2062-2078).  Contents:

  csrc/                       CUDA kernels (sm_100a) and the C-ABI, built into
  libembed_truncate_norm.so   the product; declared in include/embed_truncate_norm.h
  embedgpu.py                 host-side mirror of the proposed Go package
                              core/internal/embedgpu (ctypes over the same C-ABI)
  go/embedgpu/                the cgo package itself (source only: no Go toolchain here)

There is no CPU fallback for the computation: without the CUDA library, or without a
device, the entry points raise.
"""
__version__ = "0.1.0"
