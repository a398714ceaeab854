// LkCuda.kt — shared plumbing of the Kotlin/Native binding over include/llamakotlin_cuda.h (cinterop package
// ai.solace.llamakotlin.cuda, kotlin/llamakotlin_cuda.def).  UNCOMPILED in this repository (no Konan here); it is the
// file a maintainer of llama.kotlin adds under src/nativeMain/kotlin/ai/solace/llamakotlin/core/.
package ai.solace.llamakotlin.core

import ai.solace.llamakotlin.cuda.*
import kotlinx.cinterop.*

/** One lk_ctx per device; calls on it are serialised by the caller like everything else in the reference
 *  (single-threaded: core/GGMLScheduler.kt:165-171).  There is no CPU fallback: without a B200 this throws. */
class LkContext(val device: Int = 0) {
    val handle: CPointer<lk_ctx> = memScoped {
        val out = alloc<CPointerVar<lk_ctx>>()
        val st = lk_ctx_create(device, out.ptr)
        check(st == LK_OK.toInt()) { lk_last_error(null)?.toKString() ?: "lk_ctx_create failed: $st" }
        out.value!!
    }
    fun close() { lk_ctx_destroy(handle) }
}

/** status -> the exception class the reference throws at the cited line (include/llamakotlin_cuda.h, status codes). */
internal fun rethrow(ctx: CPointer<lk_ctx>?, st: Int) {
    if (st == LK_OK.toInt()) return
    val msg = lk_last_error(ctx)?.toKString() ?: "status $st"
    when (st) {
        LK_EINVAL_DIM.toInt(), LK_EINVAL_TYPE.toInt(), LK_EINVAL_ARG.toInt() -> throw IllegalArgumentException(msg) // GGMLComputeOps.kt:1440, :1444-1449
        LK_EBOUNDS.toInt() -> throw IndexOutOfBoundsException(msg)                                                  // GGMLTypes.kt:556-558
        LK_ENOTIMPL.toInt() -> throw NotImplementedError(msg)                                                       // GGMLComputeOps.kt:1563, :2548
        else -> throw IllegalStateException(msg)                                                                    // LK_ECUDA / LK_ENCCL / LK_ENOMEM
    }
}

/** GGMLType -> the ABI's type id (core/GGMLTypes.kt:145-153 ordinals; BITNET_1_58 has no fromValue id in the reference). */
internal fun lkTypeOf(t: GGMLType): Int = when (t) {
    GGMLType.F32 -> LK_TYPE_F32.toInt(); GGMLType.F16 -> LK_TYPE_F16.toInt()
    GGMLType.Q4_0 -> LK_TYPE_Q4_0.toInt(); GGMLType.Q4_1 -> LK_TYPE_Q4_1.toInt(); GGMLType.Q8_0 -> LK_TYPE_Q8_0.toInt()
    GGMLType.Q2_K -> LK_TYPE_Q2_K.toInt(); GGMLType.Q4_K -> LK_TYPE_Q4_K.toInt(); GGMLType.Q8_K -> LK_TYPE_Q8_K.toInt()
    GGMLType.I32 -> LK_TYPE_I32.toInt(); GGMLType.BITNET_1_58 -> LK_TYPE_BITNET_1_58.toInt()
    else -> throw NotImplementedError("type $t has no device path")
}

/** Fills an lk_tensor from a GGMLTensor: (buffer, dataOffset), never tensor.data (SURVEY F8).  `pin` is the pinned
 *  ByteArray of a host-resident tensor (pinned only for the duration of the call: GGMLAlloc.kt:361-394 may replace
 *  the array between calls, and the library never keeps a host pointer). */
internal fun fillTensor(d: lk_tensor, ga: GGMLGraphAllocator, t: GGMLTensor, pin: Pinned<ByteArray>?) {
    d.type = lkTypeOf(t.type)
    for (k in 0 until 4) { d.ne[k] = t.ne[k]; d.nb[k] = if (t.type == GGMLType.F32) t.nb[k] else 0u }
    d.data_offset = t.dataOffset
    val dev = t.buffer as? GGMLCudaBuffer
    if (dev != null) {
        d.dev_buf = dev.handle; d.host_base = null; d.buffer_size = dev.getSize()
    } else {
        val arr = ga.buffers[t.bufferId] ?: throw IllegalStateException("Tensor buffer not found for bufferId ${t.bufferId} for tensor '${t.name}'.")
        d.dev_buf = null; d.host_base = pin!!.addressOf(0); d.buffer_size = arr.size.toULong()
    }
}
