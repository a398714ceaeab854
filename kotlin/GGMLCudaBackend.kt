// GGMLCudaBackend.kt — the graph-level plug-in seam: interface GGMLBackend (core/GGMLBackend.kt:90-157),
// GGMLBackendBuffer (:45-75), GGMLBackendBufferType (:24-39) over the C ABI.  UNCOMPILED here (see llamakotlin_cuda.def).
// graphCompute hands the node list to lk_graph_compute: the fusion / chaining plan (RMS_NORM prologue, residual
// epilogue, stacked Q/K/V and gate/up, SwiGLU epilogue, chained dependent launches, the attention step) lives in the
// library (csrc/lk_planner.cu), so this file does not re-implement it.
package ai.solace.llamakotlin.core

import ai.solace.llamakotlin.cuda.*
import kotlinx.cinterop.*

class GGMLCudaBufferType(private val backend: GGMLCudaBackend) : GGMLBackendBufferType {
    override fun getName() = "CUDA"
    override fun allocBuffer(size: ULong): GGMLBackendBuffer? = GGMLCudaBuffer(backend, size)
    override fun getAlignment() = 256u
    override fun getMaxSize() = 180uL shl 30
    override fun isHost() = false        // GGMLGraphAllocator then stores buffers.add(null), GGMLAlloc.kt:300-304
}

class GGMLCudaBuffer(private val backend: GGMLCudaBackend, private val size: ULong) : GGMLBackendBuffer {
    val handle: CPointer<lk_buf> = memScoped {
        val out = alloc<CPointerVar<lk_buf>>(); rethrow(backend.lk.handle, lk_buf_alloc(backend.lk.handle, size, out.ptr)); out.value!!
    }
    override fun getType() = backend.getDefaultBufferType()
    override fun getName() = "CUDA"
    override fun getBase(): Any? = lk_buf_device_ptr(handle)
    override fun getSize() = size
    override fun free() { lk_buf_free(handle) }
    override fun initTensor(tensor: GGMLTensor) { tensor.buffer = this }
    override fun setTensor(tensor: GGMLTensor, data: ByteArray, offset: ULong, size: ULong) =
        data.usePinned { rethrow(backend.lk.handle, lk_buf_upload(handle, tensor.dataOffset + offset, it.addressOf(0), size)) }
    override fun getTensor(tensor: GGMLTensor, data: ByteArray, offset: ULong, size: ULong) =
        data.usePinned { rethrow(backend.lk.handle, lk_buf_download(handle, tensor.dataOffset + offset, it.addressOf(0), size)) }
    override fun copyTensor(src: GGMLTensor, dst: GGMLTensor): Boolean {
        val s = src.buffer as? GGMLCudaBuffer ?: return false
        rethrow(backend.lk.handle, lk_buf_copy(handle, dst.dataOffset, s.handle, src.dataOffset, calculateTensorByteSize(src)))
        return true
    }
    override fun clear(value: UByte) { rethrow(backend.lk.handle, lk_buf_clear(handle, value)) }
}

class GGMLCudaBackend(device: Int = 0) : GGMLBackend {
    val lk = LkContext(device)
    private val bufferType = GGMLCudaBufferType(this)
    private val captured = HashMap<GGMLCGraph, CPointer<lk_graph>>()      // replay path: one map lookup + lk_graph_launch

    override fun getGuid() = "llamakotlin-cuda-sm100a"
    override fun getName() = "CUDA"
    override fun getDefaultBufferType() = bufferType
    override fun allocBuffer(size: ULong) = bufferType.allocBuffer(size)
    override fun supportsBufferType(bufferType: GGMLBackendBufferType) = bufferType is GGMLCudaBufferType
    override fun supportsOp(tensor: GGMLTensor) = opOf(tensor.op) != null
    override fun offloadOp(tensor: GGMLTensor) = tensor.op == GGMLOp.MUL_MAT
    override fun synchronize() { rethrow(lk.handle, lk_sync(lk.handle)) }       // also reports a device-side barrier time-out
    override fun free() { captured.values.forEach { lk_graph_destroy(it) }; captured.clear(); lk.close() }

    private fun opOf(op: GGMLOp): Int? = when (op) {
        GGMLOp.NONE -> LK_OP_NONE.toInt(); GGMLOp.ADD -> LK_OP_ADD.toInt(); GGMLOp.MUL -> LK_OP_MUL.toInt()
        GGMLOp.MUL_MAT -> LK_OP_MUL_MAT.toInt(); GGMLOp.RMS_NORM -> LK_OP_RMS_NORM.toInt(); GGMLOp.SILU -> LK_OP_SILU.toInt()
        GGMLOp.SOFT_MAX -> LK_OP_SOFT_MAX.toInt(); GGMLOp.SCALE -> LK_OP_SCALE.toInt(); GGMLOp.ROPE -> LK_OP_ROPE.toInt()
        GGMLOp.CPY -> LK_OP_CPY.toInt(); GGMLOp.FLASH_ATTN -> LK_OP_FLASH_ATTN.toInt()
        else -> null
    }

    /** graph.nodes[0 until nNodes] (the walk of GGMLComputeOps.computeGraph, core/GGMLComputeOps.kt:2515-2523) as lk_graph_node[] */
    private fun MemScope.nodeArray(ga: GGMLGraphAllocator, graph: GGMLCGraph): CPointer<lk_graph_node> {
        val arr = allocArray<lk_graph_node>(graph.nNodes)
        for (i in 0 until graph.nNodes) {
            val nd = graph.nodes[i]!!
            arr[i].op = opOf(nd.op) ?: throw NotImplementedError("Operation ${nd.op} not implemented")       // :2548
            val srcs = nd.src.takeWhile { it != null }.take(4)
            arr[i].n_src = srcs.size
            fillTensor(arr[i].dst, ga, nd, null)              // device-resident graphs only: see supportsBufferType
            srcs.forEachIndexed { k, s -> fillTensor(arr[i].src[k], ga, s!!, null) }
            for (k in 0 until 4) arr[i].op_params[k] = nd.opParams[k]
        }
        return arr
    }

    /** core/GGMLBackend.kt:145; exceptions become FAILED like GGMLCpuBackend.kt:168-175.  A graph seen before is one
     *  CUDA-graph launch.  A graph whose node list or offsets change gets a new capture: drop the old one with forget(). */
    override fun graphCompute(graph: GGMLCGraph): GGMLStatus = try {
        val g = captured.getOrPut(graph) {
            memScoped {
                val nodes = nodeArray(graph.allocator!!, graph)
                rethrow(lk.handle, lk_capture_begin(lk.handle))
                val st = lk_graph_compute(lk.handle, nodes, graph.nNodes, 0)
                val out = alloc<CPointerVar<lk_graph>>()
                val st2 = lk_capture_end(lk.handle, out.ptr)
                rethrow(lk.handle, st); rethrow(lk.handle, st2)
                out.value!!
            }
        }
        rethrow(lk.handle, lk_graph_launch(lk.handle, g))
        GGMLStatus.SUCCESS
    } catch (e: Exception) {
        println("CUDA backend graph computation failed: ${e.message}")
        GGMLStatus.FAILED
    }

    fun forget(graph: GGMLCGraph) { captured.remove(graph)?.let { lk_graph_destroy(it) }; capturedIo.remove(graph)?.let { lk_graph_destroy(it) } }

    /** One host copy of a step: `pinned` (from lk_host_alloc) <-> the bytes of `tensor` in its device buffer. */
    class HostCopy(val upload: Boolean, val tensor: GGMLTensor, val pinned: COpaquePointer, val bytes: ULong)

    private val capturedIo = HashMap<GGMLCGraph, CPointer<lk_graph>>()

    /** setTensor -> graphCompute -> getTensor as ONE submission (lk_graph_compute_io): every upload is issued right
     *  before the first launch that reads its region, every download right after the last launch that writes it, on copy
     *  streams beside the compute stream.  Split the copies by consumer (e.g. per group of layers); the downloads are
     *  complete after synchronize().  The copies are nodes of the captured graph: the same pinned addresses every step. */
    fun graphCompute(graph: GGMLCGraph, copies: List<HostCopy>): GGMLStatus = try {
        val g = capturedIo.getOrPut(graph) {
            memScoped {
                val nodes = nodeArray(graph.allocator!!, graph)
                val io = allocArray<lk_host_io>(copies.size)
                copies.forEachIndexed { k, c ->
                    io[k].direction = if (c.upload) LK_IO_UPLOAD.toInt() else LK_IO_DOWNLOAD.toInt()
                    io[k].buf = (c.tensor.buffer as GGMLCudaBuffer).handle
                    io[k].offset = c.tensor.dataOffset
                    io[k].bytes = c.bytes
                    io[k].host = c.pinned
                }
                rethrow(lk.handle, lk_capture_begin(lk.handle))
                val st = lk_graph_compute_io(lk.handle, nodes, graph.nNodes, 0, io, copies.size)
                val out = alloc<CPointerVar<lk_graph>>()
                val st2 = lk_capture_end(lk.handle, out.ptr)
                rethrow(lk.handle, st); rethrow(lk.handle, st2)
                out.value!!
            }
        }
        rethrow(lk.handle, lk_graph_launch(lk.handle, g))
        GGMLStatus.SUCCESS
    } catch (e: Exception) {
        println("CUDA backend graph computation failed: ${e.message}")
        GGMLStatus.FAILED
    }
}

/** Registration next to the CPU and Metal entries (core/GGMLBackend.kt:162-181):
 *    GGMLBackendRegistry.register(GGMLBackendRegistration("CUDA", { _ -> GGMLCudaBackend(0) }, null))
 *    val graph = createGraph(size, GGMLCudaBackend(0))        // GGMLGraph.kt:2596-2601 -> GGMLGraphAllocator(backend)  */
