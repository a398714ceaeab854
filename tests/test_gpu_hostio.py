"""Host copies inside a step: lk_buf_upload_async / lk_buf_download_async and lk_graph_compute_io
(GGMLCudaBackend.graphCompute(graph, hostIO=...)).  The reference's sequence is setTensor -> graphCompute -> getTensor
(core/GGMLBackend.kt:95-110, :145); the pipelined form must deliver the same bytes: it is compared bit for bit with that
sequence, as a captured replay (the copies are nodes of the CUDA graph and read the host arrays at every replay) and
issued directly."""
import ctypes as C

import numpy as np
import pytest

import llama_kotlin_b200 as lk
from conftest import same_f32_bits, synth_acts, synth_weights
from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType
from llama_kotlin_b200 import _binding as B
from test_gpu_parity import QT

pytestmark = pytest.mark.gpu


def pinned(be, nbytes):
    p = C.c_void_p()
    B.check(B.lib().lk_host_alloc(be.lk.handle, nbytes, C.byref(p)), be.lk.handle)
    return np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(p.value))


def _independent_gemvs(O, be, layers, M, K, tname="Q4_0"):
    """`layers` GEMVs, each with its own weights, x and dst (the shape of bench.py's step), x and dst contiguous per layer."""
    gt, ot = QT[tname]
    ga = lk.GGMLGraphAllocator(be, size=layers * (M * K) + layers * (M + K) * 8 + (1 << 20))
    acts = ga.allocateTensor(GGMLTensor(GGMLType.F32, [layers * K]))
    dsts = ga.allocateTensor(GGMLTensor(GGMLType.F32, [layers * M]))
    nodes = []
    for l in range(layers):
        a = ga.allocateTensor(GGMLTensor(gt, [K, M]))
        ga.setBytes(a, O.quantize(ot, synth_weights(8100 + l, M, K)))
        b = GGMLTensor(GGMLType.F32, [1, K], bufferId=0, dataOffset=acts.dataOffset + l * K * 4, buffer=ga.deviceBuffer)
        nodes.append(GGMLTensor(GGMLType.F32, [1, M], op=GGMLOp.MUL_MAT, src=[a, b] + [None] * 8, bufferId=0,
                                dataOffset=dsts.dataOffset + l * M * 4, buffer=ga.deviceBuffer))
    return ga, acts, dsts, nodes


def _views(ga, acts, dsts, layers, M, K, xh, yh):
    io = []
    for l in range(layers):
        xv = GGMLTensor(GGMLType.F32, [K], bufferId=0, dataOffset=acts.dataOffset + l * K * 4, buffer=ga.deviceBuffer)
        yv = GGMLTensor(GGMLType.F32, [M], bufferId=0, dataOffset=dsts.dataOffset + l * M * 4, buffer=ga.deviceBuffer)
        io.append(("in", xv, xh[l * K:(l + 1) * K]))
        io.append(("out", yv, yh[l * M:(l + 1) * M]))
    return io


@pytest.mark.parametrize("use_graphs", [True, False])
@pytest.mark.parametrize("M,K", [(256, 512), (4096, 4096)])
def test_pipelined_copies_deliver_what_the_serial_sequence_does(O, use_graphs, M, K):
    layers = 6
    be = lk.GGMLCudaBackend(0)
    be.useCudaGraphs = use_graphs
    ga, acts, dsts, nodes = _independent_gemvs(O, be, layers, M, K)
    xh = pinned(be, layers * K * 4).view(np.float32)
    yh = pinned(be, layers * M * 4).view(np.float32)
    serial = np.empty(layers * M, np.float32)
    g_serial, g_io = GGMLCGraph(nodes=nodes, allocator=ga), GGMLCGraph(nodes=nodes, allocator=ga)
    io = _views(ga, acts, dsts, layers, M, K, xh, yh)
    for rep in range(3):                      # rep > 0: a replay of the capture, with NEW host contents each time
        xh[:] = np.concatenate([synth_acts(8200 + 10 * rep + l, K) for l in range(layers)])
        ga.deviceBuffer.setTensor(acts, xh, 0, xh.nbytes)
        assert be.graphCompute(g_serial) == lk.GGMLStatus.SUCCESS
        ga.deviceBuffer.getTensor(dsts, serial.view(np.uint8), 0, serial.nbytes)
        ga.setBytes(acts, np.zeros(layers * K, np.float32))      # the pipelined step must bring its own activations
        ga.setBytes(dsts, np.full(layers * M, np.nan, np.float32))
        yh[:] = -1.0
        assert be.graphCompute(g_io, hostIO=io) == lk.GGMLStatus.SUCCESS
        be.synchronize()
        assert float(np.max(np.abs(serial))) > 0
        assert same_f32_bits(yh, serial), rep
    be.free()


def test_pipelined_copies_around_a_fused_dependent_graph(O):
    """RMS_NORM -> MUL_MAT -> ADD (folded into one launch by the planner): the upload feeds the folded RMS_NORM, the download
    follows the folded ADD."""
    hidden = 512
    gt, ot = QT["Q4_0"]
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=hidden * hidden + 64 * hidden * 4 + (1 << 20))

    def vec(shape, data=None):
        t = ga.allocateTensor(GGMLTensor(GGMLType.F32, shape))
        if data is not None:
            ga.setBytes(t, data)
        return t
    W = ga.allocateTensor(GGMLTensor(gt, [hidden, hidden]))
    ga.setBytes(W, O.quantize(ot, synth_weights(8300, hidden, hidden)))
    x, gain, res = vec([hidden]), vec([hidden], 1.0 + 0.1 * synth_acts(8301, hidden)), vec([hidden])
    xn, y, out = vec([hidden]), vec([1, hidden]), vec([hidden])

    def node(op, t, src, eps=None):
        n = GGMLTensor(GGMLType.F32, [hidden], op=op, src=list(src) + [None] * (10 - len(src)), bufferId=0, dataOffset=t.dataOffset,
                       buffer=ga.deviceBuffer)
        if eps is not None:
            n.opParams[0] = int(np.float32(eps).view(np.int32))
        return n
    n_rms = node(GGMLOp.RMS_NORM, xn, [x, gain], eps=1e-6)
    xn_col = GGMLTensor(GGMLType.F32, [1, hidden], bufferId=0, dataOffset=xn.dataOffset, buffer=ga.deviceBuffer)
    n_mm = GGMLTensor(GGMLType.F32, [1, hidden], op=GGMLOp.MUL_MAT, src=[W, xn_col] + [None] * 8, bufferId=0, dataOffset=y.dataOffset,
                      buffer=ga.deviceBuffer)
    y_flat = GGMLTensor(GGMLType.F32, [hidden], bufferId=0, dataOffset=y.dataOffset, buffer=ga.deviceBuffer)
    n_add = node(GGMLOp.ADD, out, [y_flat, res])
    nodes = [n_rms, n_mm, n_add]
    xh = pinned(be, 2 * hidden * 4).view(np.float32)
    oh = pinned(be, hidden * 4).view(np.float32)
    io = [("in", x, xh[:hidden]), ("in", res, xh[hidden:]), ("out", out, oh)]
    g_serial, g_io = GGMLCGraph(nodes=nodes, allocator=ga), GGMLCGraph(nodes=nodes, allocator=ga)
    for rep in range(2):
        xh[:hidden] = synth_acts(8310 + rep, hidden)
        xh[hidden:] = synth_acts(8320 + rep, hidden)
        ga.setBytes(x, xh[:hidden])
        ga.setBytes(res, xh[hidden:])
        assert be.graphCompute(g_serial) == lk.GGMLStatus.SUCCESS
        want = ga.getFloats(out).copy()
        ga.setBytes(x, np.zeros(hidden, np.float32))
        ga.setBytes(res, np.zeros(hidden, np.float32))
        ga.setBytes(out, np.full(hidden, np.nan, np.float32))
        assert be.graphCompute(g_io, hostIO=io) == lk.GGMLStatus.SUCCESS
        be.synchronize()
        assert float(np.max(np.abs(want))) > 0 and same_f32_bits(oh, want), rep
    be.free()


def test_async_copies_issued_directly(O):
    """setTensorAsync -> computeMatMul -> getTensorAsync -> synchronize, twice with different data."""
    M, K = 320, 256
    gt, ot = QT["Q8_0"]
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=M * K * 2 + (M + K) * 16 + (1 << 20))
    a = ga.allocateTensor(GGMLTensor(gt, [K, M]))
    blocks = O.quantize(ot, synth_weights(8400, M, K))
    ga.setBytes(a, blocks)
    b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, K]))
    d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, M]))
    xh = pinned(be, K * 4).view(np.float32)
    yh = pinned(be, M * 4).view(np.float32)
    for rep in range(2):
        xh[:] = synth_acts(8410 + rep, K)
        yh[:] = np.nan
        ga.deviceBuffer.setTensorAsync(b, xh, 0, xh.nbytes)
        lk.computeMatMul(ga, None, a, b, d)
        ga.deviceBuffer.getTensorAsync(d, yh, 0, yh.nbytes)
        be.synchronize()
        ga.setBytes(b, xh)
        lk.computeMatMul(ga, None, a, b, d)
        assert same_f32_bits(yh, ga.getFloats(d)), rep
    be.free()


def test_graph_compute_io_rejects_a_copy_without_an_address():
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=1 << 20)
    io = (B.lk_host_io * 1)()
    io[0].direction = B.LK_IO_UPLOAD
    io[0].buf = ga.deviceBuffer.handle
    io[0].offset, io[0].bytes, io[0].host = 0, 64, None
    st = B.lib().lk_graph_compute_io(be.lk.handle, None, 0, 0, io, 1)
    assert st == B.LK_EINVAL_ARG and b"host address" in B.lib().lk_last_error(be.lk.handle)
    io[0].host = 1 << 20
    io[0].offset = 2 << 20           # past the end of the buffer
    host = np.zeros(64, np.uint8)
    io[0].host = host.ctypes.data
    st = B.lib().lk_graph_compute_io(be.lk.handle, None, 0, 0, io, 1)
    assert st != B.LK_OK
    be.free()
