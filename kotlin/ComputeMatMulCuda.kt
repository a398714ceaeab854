// ComputeMatMulCuda.kt — op-level drop-ins: the same signatures and the same exceptions at the same conditions as
// computeMatMul / computeAdd / computeMul / dequantizeTensor / quantizeTensor (core/GGMLComputeOps.kt:1435, :735, :811,
// :918, :1040) and the three node ops the reference declares but leaves empty (RMS_NORM, SILU, SOFT_MAX:
// :2633-2645, :2548).  UNCOMPILED here (see kotlin/llamakotlin_cuda.def).
// Call sites to switch: GGMLComputeOps.computeMulMat (:2647-2652) and computeMatMulRet (:905-916).
package ai.solace.llamakotlin.core

import ai.solace.llamakotlin.cuda.*
import kotlinx.cinterop.*

private val defaultCtx: LkContext by lazy { LkContext(0) }
private fun ctxOf(ga: GGMLGraphAllocator): CPointer<lk_ctx> = (ga.backend as? GGMLCudaBackend)?.lk?.handle ?: defaultCtx.handle

/** Runs `call` with lk_tensor descriptors of `ts`, host-resident ones pinned for exactly the duration of the call. */
private inline fun withTensors(ga: GGMLGraphAllocator, vararg ts: GGMLTensor?, call: (List<CPointer<lk_tensor>?>) -> Int) {
    val pins = ts.map { t -> if (t == null || t.buffer is GGMLCudaBuffer) null else (ga.buffers[t.bufferId]
        ?: throw IllegalStateException("Tensor buffer not found for bufferId ${t.bufferId}")).pin() }
    try {
        memScoped {
            val ds = ts.mapIndexed { i, t -> t?.let { alloc<lk_tensor>().also { d -> fillTensor(d, ga, it, pins[i]) }.ptr } }
            rethrow(ctxOf(ga), call(ds))
        }   // host tensors: dst is valid here (the call staged H2D, ran, copied back and synchronised)
    } finally { pins.forEach { it?.unpin() } }
}

fun computeMatMulCuda(graphAllocator: GGMLGraphAllocator, @Suppress("unused") context: GGMLContext,
                      a: GGMLTensor, b: GGMLTensor, dst: GGMLTensor) =
    withTensors(graphAllocator, a, b, dst) { d -> lk_compute_matmul(ctxOf(graphAllocator), d[0], d[1], d[2]) }

fun computeAddCuda(graphAllocator: GGMLGraphAllocator, @Suppress("unused") context: GGMLContext, a: GGMLTensor, b: GGMLTensor, dst: GGMLTensor) =
    withTensors(graphAllocator, a, b, dst) { d -> lk_compute_add(ctxOf(graphAllocator), d[0], d[1], d[2]) }

fun computeMulCuda(graphAllocator: GGMLGraphAllocator, @Suppress("unused") context: GGMLContext, a: GGMLTensor, b: GGMLTensor, dst: GGMLTensor) =
    withTensors(graphAllocator, a, b, dst) { d -> lk_compute_mul(ctxOf(graphAllocator), d[0], d[1], d[2]) }

fun computeSiluCuda(graphAllocator: GGMLGraphAllocator, x: GGMLTensor, dst: GGMLTensor) =
    withTensors(graphAllocator, x, dst) { d -> lk_compute_silu(ctxOf(graphAllocator), d[0], d[1]) }

fun computeRmsNormCuda(graphAllocator: GGMLGraphAllocator, x: GGMLTensor, weight: GGMLTensor?, eps: Float, dst: GGMLTensor) =
    withTensors(graphAllocator, x, weight, dst) { d -> lk_compute_rms_norm(ctxOf(graphAllocator), d[0], d[1], eps, d[2]) }

fun computeSoftMaxCuda(graphAllocator: GGMLGraphAllocator, x: GGMLTensor, dst: GGMLTensor) =
    withTensors(graphAllocator, x, dst) { d -> lk_compute_softmax(ctxOf(graphAllocator), d[0], d[1]) }

/** dequantizeTensor: the result lives in result.data (FloatArray), like the reference (:918-1038). */
fun dequantizeTensorCuda(graphAllocator: GGMLGraphAllocator, tensor: GGMLTensor): GGMLTensor {
    val n = tensor.numElements().toInt()
    val out = FloatArray(n)
    val res = GGMLTensor(GGMLType.F32).apply { ne = tensor.ne.copyOf(); data = out }
    val pin = if (tensor.buffer is GGMLCudaBuffer) null else graphAllocator.buffers[tensor.bufferId]!!.pin()
    try {
        out.usePinned { po ->
            memScoped {
                val src = alloc<lk_tensor>().also { fillTensor(it, graphAllocator, tensor, pin) }
                val dst = alloc<lk_tensor>().apply {
                    type = LK_TYPE_F32.toInt(); for (k in 0 until 4) { ne[k] = tensor.ne[k]; nb[k] = 0u }
                    host_base = po.addressOf(0); buffer_size = (n * 4).toULong(); data_offset = 0u
                }
                rethrow(ctxOf(graphAllocator), lk_compute_dequantize(ctxOf(graphAllocator), src.ptr, dst.ptr))
            }
        }
    } finally { pin?.unpin() }
    return res
}

/** quantizeTensor (:1040-1433) for Q8_0 / Q4_0 / Q4_1: bytes in result.data, bufferId = -1 like the reference (SURVEY F8). */
fun quantizeTensorCuda(graphAllocator: GGMLGraphAllocator, tensor: GGMLTensor, targetType: GGMLType): GGMLTensor {
    val n = tensor.numElements()
    val out = ByteArray((n / 32 * targetType.byteSize.toLong()).toInt())
    val res = GGMLTensor(targetType).apply { ne = tensor.ne.copyOf(); data = out }
    val pin = if (tensor.buffer is GGMLCudaBuffer) null else graphAllocator.buffers[tensor.bufferId]!!.pin()
    try {
        out.usePinned { po ->
            memScoped {
                val src = alloc<lk_tensor>().also { fillTensor(it, graphAllocator, tensor, pin) }
                val dst = alloc<lk_tensor>().apply {
                    type = lkTypeOf(targetType); for (k in 0 until 4) { ne[k] = tensor.ne[k]; nb[k] = 0u }
                    host_base = po.addressOf(0); buffer_size = out.size.toULong(); data_offset = 0u
                }
                rethrow(ctxOf(graphAllocator), lk_compute_quantize(ctxOf(graphAllocator), src.ptr, dst.ptr))
            }
        }
    } finally { pin?.unpin() }
    return res
}
