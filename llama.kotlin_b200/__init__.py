"""llama.kotlin quantized-matmul hot path on B200 (sm_100a).

Product = libllamakotlin_cuda.so behind include/llamakotlin_cuda.h; this package is
the host-side mirror of the reference's operator / backend interface over that C ABI.
No CPU fallback exists anywhere in this package."""
from ._binding import (IllegalArgumentException, IllegalStateException,  # noqa: F401
                       IndexOutOfBoundsException)
from .ggml_alloc import GGMLGraphAllocator  # noqa: F401
from .ggml_compute_ops import (LkContext, computeAdd, computeMatMul, computeMatMulFused, computeMul,  # noqa: F401
                               computeRmsNorm, computeSilu, computeSiluMul, computeSoftMax,
                               KVCache, computeAttentionDecode, computeAttentionStep, computeKvAppend, computeRope, computeScale, ropeTable,
                               default_context, dequantizeTensor, computeQuantize, quantizeTensor, computeMatMulChain)
from .ggml_cuda_backend import (GGMLCudaBackend, GGMLCudaBuffer, GGMLCudaBufferType)  # noqa: F401
from .llama_layer import (DecoderLayer, alloc_layer, append_layer_nodes, build_decoder_graph,  # noqa: F401
                          weight_shapes)
from .sharded import ShardGroup, shard_byte_range, shard_rows  # noqa: F401
from .ggml_types import (GGMLCGraph, GGMLContext, GGMLOp, GGMLStatus, GGMLTensor,  # noqa: F401
                         GGMLType, calculateContiguousStrides)
