// lk_planner.cu — graph level of the C ABI: GGMLBackend.graphCompute (core/GGMLBackend.kt:145) as a library call.
//
// The reference walks graph.nodes[0 .. nNodes) with a `when (node.op)` (GGMLComputeOps.computeGraph / computeNode,
// core/GGMLComputeOps.kt:2515-2550).  lk_graph_compute does the same walk, but first plans which neighbours fold
// into the decode GEMV launches (op order of LlamaDecoderLayer.forward, model/LlamaModel.kt:253-285):
//   RMS_NORM -> MUL_MAT (x up to 3 on the same activations), MUL_MAT -> ADD, MUL_MAT, MUL_MAT -> SILU -> MUL,
//   runs of dependent launches as one persistent launch, ROPE, ROPE, CPY, CPY, FLASH_ATTN as one attention step.
// Every folded node's dst is still written, so the graph's intermediates stay valid whatever reads them later.
// lk_graph_plan is the planning half on its own: pure host logic, no device needed (tests/test_planner_cpu.py).
// The Python mirror llama.kotlin_b200/ggml_cuda_backend.py::_plan / _runNodes is the same algorithm; a Kotlin
// maintainer binds these two entry points instead of porting it (kotlin/GGMLCudaBackend.kt).
#include <string.h>

#include <algorithm>
#include <map>
#include <set>
#include <tuple>
#include <vector>

#include "lk_shim.h"

namespace {

typedef std::tuple<const void *, uint64_t, int64_t> Key;      // (buffer, dataOffset, elements): "the same tensor"
struct Span { const void *buf; uint64_t lo, hi; };            // bytes a tensor occupies in its buffer

int64_t nelem(const lk_tensor &t) {
    int64_t n = 1;
    for (int d = 0; d < LK_MAX_DIMS; d++) n *= (t.ne[d] > 0 ? t.ne[d] : (t.ne[d] == 0 && d == 0 ? 0 : 1));
    return n;
}
const void *buf_of(const lk_tensor &t) { return t.dev_buf ? (const void *)t.dev_buf : (const void *)t.host_base; }
Key key_of(const lk_tensor &t) { return Key(buf_of(t), t.data_offset, nelem(t)); }
Span span_of(const lk_tensor &t) {
    const int64_t nb = lk_tensor_nbytes(&t);
    return Span{buf_of(t), t.data_offset, t.data_offset + (uint64_t)(nb > 0 ? nb : 0)};
}
bool overlaps(const Span &a, const Span &b) { return a.buf == b.buf && a.lo < b.hi && b.lo < a.hi; }
bool same_span(const Span &a, const Span &b) { return a.buf == b.buf && a.lo == b.lo && a.hi == b.hi; }
bool is_q(int t) { return t == LK_TYPE_Q4_0 || t == LK_TYPE_Q4_1 || t == LK_TYPE_Q8_0; }

struct Planner {
    const lk_graph_node *nodes;
    int n;
    std::map<Key, std::vector<std::pair<int, int>>> consumers;   // tensor -> (node, source slot) that read it
    std::map<int, int> rms_of, add_of;                           // GEMV node -> folded RMS_NORM / ADD node
    std::map<int, std::vector<int>> groups;                      // leader -> stacked GEMV nodes (leader first)
    std::map<int, std::pair<int, int>> pairs;                    // leader of a gate/up pair -> (SILU, MUL) nodes
    std::map<int, std::vector<int>> steps;                       // ROPE(q) node -> the five nodes of an attention step
    std::set<int> skip, gemv;

    bool is_gemv(int i) const {
        const lk_graph_node &t = nodes[i];
        return t.op == LK_OP_MUL_MAT && t.n_src >= 2 && is_q(t.src[0].type) && t.src[1].ne[0] == 1 && t.dst.type == LK_TYPE_F32;
    }
    std::vector<std::pair<int, int>> cons_after(const lk_tensor &t, int i) const {
        std::vector<std::pair<int, int>> out;
        auto it = consumers.find(key_of(t));
        if (it != consumers.end())
            for (const auto &c : it->second) if (c.first > i) out.push_back(c);
        return out;
    }

    void plan() {
        for (int j = 0; j < n; j++)
            for (int s = 0; s < nodes[j].n_src && s < 4; s++) consumers[key_of(nodes[j].src[s])].push_back({j, s});
        // RMS_NORM folds into the GEMVs that consume it
        for (int i = 0; i < n; i++) {
            const lk_graph_node &t = nodes[i];
            if (t.op != LK_OP_RMS_NORM || t.dst.type != LK_TYPE_F32 || nelem(t.dst) != t.dst.ne[0]) continue;
            const auto cons = cons_after(t.dst, i);
            if (cons.empty()) continue;
            bool all = true;
            int last = 0;
            for (const auto &c : cons) { all = all && is_gemv(c.first) && c.second == 1; last = std::max(last, c.first); }
            if (!all) continue;
            // a later writer of the same storage before the last consumer would change what they read
            bool clean = true;
            for (int k = i + 1; k <= last && clean; k++) {
                if (nodes[k].op == LK_OP_NONE) continue;
                const Span sk = span_of(nodes[k].dst);
                clean = !overlaps(sk, span_of(t.dst)) && !overlaps(sk, span_of(t.src[0])) &&
                        (t.n_src < 2 || !overlaps(sk, span_of(t.src[1])));
            }
            if (!clean) continue;
            for (const auto &c : cons) rms_of[c.first] = i;
            skip.insert(i);
        }
        // ADD folds into the GEMV that feeds it
        for (int i = 0; i < n; i++) {
            if (!is_gemv(i)) continue;
            const auto cons = cons_after(nodes[i].dst, i);
            if (cons.size() != 1) continue;
            const int j = cons[0].first, sidx = cons[0].second;
            const lk_graph_node &a = nodes[j];
            if (a.op != LK_OP_ADD || a.dst.type != LK_TYPE_F32 || skip.count(j) || a.n_src < 2 || sidx > 1) continue;
            const lk_tensor &res = a.src[1 - sidx];
            if (nelem(res) != nelem(nodes[i].dst)) continue;
            // the ADD moves up to the matmul's position: nothing in between may write its operand or touch its dst
            bool ok = true;
            for (int k = i; k < j && ok; k++) {
                if (nodes[k].op != LK_OP_NONE)
                    ok = !overlaps(span_of(nodes[k].dst), span_of(res)) && !overlaps(span_of(nodes[k].dst), span_of(a.dst));
                for (int s = 0; s < nodes[k].n_src && s < 4 && ok; s++) ok = !overlaps(span_of(nodes[k].src[s]), span_of(a.dst));
            }
            if (!ok) continue;
            add_of[i] = j;
            skip.insert(j);
        }
        // stacked matmuls (same activations, weights of one type and shape) and the SwiGLU pair
        for (int i = 0; i < n;) {
            if (skip.count(i) || !is_gemv(i) || add_of.count(i)) { i++; continue; }
            const lk_graph_node &t = nodes[i];
            std::vector<int> members{i};
            int k = i + 1;
            while (k < n && members.size() < 3) {
                if (skip.count(k)) { k++; continue; }
                const lk_graph_node &u = nodes[k];
                bool ok = is_gemv(k) && !add_of.count(k) && key_of(u.src[1]) == key_of(t.src[1]) && u.src[0].type == t.src[0].type &&
                          memcmp(u.src[0].ne, t.src[0].ne, sizeof(t.src[0].ne)) == 0 &&
                          (rms_of.count(k) ? rms_of[k] : -1) == (rms_of.count(i) ? rms_of[i] : -1);
                for (size_t m = 0; m < members.size() && ok; m++) ok = !overlaps(span_of(u.src[1]), span_of(nodes[members[m]].dst));
                if (!ok) break;
                members.push_back(k);
                k++;
            }
            if (members.size() > 1) {
                groups[i] = members;
                for (size_t m = 1; m < members.size(); m++) skip.insert(members[m]);
                if (members.size() == 2) {
                    const int g = members[0], u2 = members[1];
                    const auto cg = cons_after(nodes[g].dst, g), cu = cons_after(nodes[u2].dst, u2);
                    if (cg.size() == 1 && cu.size() == 1 && nodes[cg[0].first].op == LK_OP_SILU && nodes[cu[0].first].op == LK_OP_MUL) {
                        const int si = cg[0].first, mi = cu[0].first;
                        const auto cs = cons_after(nodes[si].dst, si);
                        if (si == u2 + 1 && mi == si + 1 && cs.size() == 1 && cs[0].first == mi && nodes[mi].dst.type == LK_TYPE_F32 &&
                            nelem(nodes[mi].dst) == nelem(nodes[g].dst)) {
                            pairs[i] = {si, mi};
                            skip.insert(si);
                            skip.insert(mi);
                        }
                    }
                }
            }
            i = members.back() + 1;
        }
        for (int i = 0; i < n; i++) if (is_gemv(i)) gemv.insert(i);
        plan_attention_steps();
    }

    // ROPE(q), ROPE(k), CPY(k' -> key cache), CPY(v -> value cache), FLASH_ATTN: the core of one decode step of
    // LlamaAttention.forward (model/LlamaAttention.kt:87-123) as one launch; only the exact five-node run with one token
    void plan_attention_steps() {
        for (int i = 0; i + 4 < n; i++) {
            bool any_skip = false;
            for (int k = i; k < i + 5; k++) any_skip = any_skip || skip.count(k);
            if (any_skip) continue;
            const lk_graph_node &rq = nodes[i], &rk = nodes[i + 1], &ck = nodes[i + 2], &cv = nodes[i + 3], &fa = nodes[i + 4];
            if (rq.op != LK_OP_ROPE || rk.op != LK_OP_ROPE || ck.op != LK_OP_CPY || cv.op != LK_OP_CPY || fa.op != LK_OP_FLASH_ATTN) continue;
            if (rq.n_src < 3 || rk.n_src < 3 || ck.n_src < 3 || cv.n_src < 3 || fa.n_src < 4) continue;
            const int hd = rq.op_params[0];
            const Span pos = span_of(rq.src[1]), table = span_of(rq.src[2]);
            if (!(same_span(span_of(rk.src[1]), pos) && same_span(span_of(rk.src[2]), table) && rk.op_params[0] == hd &&
                  same_span(span_of(ck.src[2]), pos) && same_span(span_of(cv.src[2]), pos) && same_span(span_of(fa.src[3]), pos)))
                continue;
            const lk_tensor &kc = ck.src[1], &vc = cv.src[1];
            if (kc.ne[0] != hd || memcmp(kc.ne, vc.ne, sizeof(kc.ne)) != 0) continue;
            const int64_t rowlen = kc.ne[0] * kc.ne[1];
            if (!(nelem(rq.dst) == rowlen && nelem(rk.dst) == rowlen && nelem(ck.src[0]) == rowlen && nelem(cv.src[0]) == rowlen &&
                  nelem(fa.dst) == rowlen))
                continue;                                   // one token only
            if (!(same_span(span_of(ck.src[0]), span_of(rk.dst)) && same_span(span_of(fa.src[0]), span_of(rq.dst)) &&
                  same_span(span_of(fa.src[1]), span_of(kc)) && same_span(span_of(fa.src[2]), span_of(vc))))
                continue;
            // the five dsts and the raw inputs must not alias each other
            const Span sp[8] = {span_of(rq.dst), span_of(rk.dst), span_of(fa.dst), span_of(kc), span_of(vc), span_of(rq.src[0]),
                                span_of(rk.src[0]), span_of(cv.src[0])};
            bool alias = false;
            for (int x = 0; x < 8 && !alias; x++)
                for (int y = x + 1; y < 8 && !alias; y++) alias = overlaps(sp[x], sp[y]);
            if (alias) continue;
            steps[i] = {i, i + 1, i + 2, i + 3, i + 4};
            for (int k = i + 1; k < i + 5; k++) skip.insert(k);
        }
    }

    // (reads, writes) of the launch made at GEMV node i, weights left out
    void link_spans(int i, std::vector<Span> &reads, std::vector<Span> &writes) const {
        const lk_graph_node &node = nodes[i];
        writes.push_back(span_of(node.dst));
        if (rms_of.count(i)) {
            const lk_graph_node &r = nodes[rms_of.at(i)];
            reads.push_back(span_of(r.src[0]));
            writes.push_back(span_of(r.dst));
        } else {
            reads.push_back(span_of(node.src[1]));
        }
        if (groups.count(i)) {
            const auto &m = groups.at(i);
            for (size_t q = 1; q < m.size(); q++) writes.push_back(span_of(nodes[m[q]].dst));
            if (pairs.count(i)) { writes.push_back(span_of(nodes[pairs.at(i).first].dst)); writes.push_back(span_of(nodes[pairs.at(i).second].dst)); }
        } else if (add_of.count(i)) {
            const lk_graph_node &a = nodes[add_of.at(i)];
            reads.push_back(span_of(key_of(a.src[0]) == key_of(node.dst) ? a.src[1] : a.src[0]));
            writes.push_back(span_of(a.dst));
        }
    }
    bool depends_on(int i, int j) const {
        std::vector<Span> ri, wi, rj, wj;
        link_spans(i, ri, wi);
        link_spans(j, rj, wj);
        for (const Span &r : ri) for (const Span &w : wj) if (overlaps(r, w)) return true;
        return false;
    }

    // the launch list: what _runNodes issues, in order
    std::vector<lk_plan_launch> launches(int chain_mode, int max_chain) {
        std::vector<lk_plan_launch> out;
        std::set<int> stored;
        std::vector<int> pending;
        int chain_id = 0;
        auto record = [&](int i, int chain) {
            lk_plan_launch L;
            memset(&L, 0xFF, sizeof(L));                 // every index field -1
            L.node = i; L.chain = chain; L.n_extra = 0; L.store_b = 0;
            if (steps.count(i)) {
                L.kind = LK_PLAN_ATTN_STEP;
                for (int q = 0; q < 5; q++) L.attn[q] = steps[i][(size_t)q];
            } else if (gemv.count(i) && (groups.count(i) || rms_of.count(i) || add_of.count(i))) {
                L.kind = LK_PLAN_FUSED;
                if (groups.count(i)) {
                    const auto &m = groups[i];
                    L.n_extra = (int)m.size() - 1;
                    for (size_t q = 1; q < m.size(); q++) L.extra[q - 1] = m[q];
                    if (pairs.count(i)) { L.silu = pairs[i].first; L.mul = pairs[i].second; }
                } else if (add_of.count(i)) {
                    L.add = add_of[i];
                }
                if (rms_of.count(i)) {
                    L.rms = rms_of[i];
                    L.store_b = stored.count(L.rms) ? 0 : 1;
                    stored.insert(L.rms);
                }
            } else {
                L.kind = LK_PLAN_NODE;
            }
            out.push_back(L);
        };
        auto flush = [&]() {
            for (size_t c = 0; c < pending.size(); c += (size_t)max_chain) {
                const size_t e = std::min(pending.size(), c + (size_t)max_chain);
                const int id = (e - c >= 2) ? chain_id++ : -1;
                for (size_t q = c; q < e; q++) record(pending[q], id);
            }
            pending.clear();
        };
        // the planner's own choice (chain_mode < 0): chain where it was measured to pay: graphs with an attention step
        // between the runs (either chained kernel), and graphs whose decode GEMVs all have K % 256 == 0 (K1m takes Q4_0 / Q4_1 and
        // short Q8_0 rows, K1c the rest; per 7B layer, chained against separate launches: Q4_0 36.5 / 39.4 us, Q4_1 36.6 / 39.2,
        // Q8_0 54.9 / 61.5; profiles/r02j_layer_ab_per_warp_prep.txt, r02m_layer_types.txt)
        bool k1m = !gemv.empty();
        for (int i : gemv) k1m = k1m && nodes[i].src[0].ne[0] % 256 == 0;      // (every decode GEMV type has a chained kernel: is_gemv)
        const bool chain = chain_mode < 0 ? (!steps.empty() || k1m) : chain_mode != 0;
        for (int i = 0; i < n; i++) {
            if (skip.count(i) || (!pending.empty() && nodes[i].op == LK_OP_NONE)) continue;
            if (chain && gemv.count(i)) {
                // only DEPENDENT launches are chained; independent GEMVs keep the pairwise-overlapped launches
                if (!pending.empty() && !depends_on(i, pending.back())) flush();
                pending.push_back(i);
                continue;
            }
            flush();
            record(i, -1);
        }
        flush();
        return out;
    }
};

}  // namespace

extern "C" {

lk_status lk_graph_plan(const lk_graph_node *nodes, int32_t n, int32_t chain_mode, lk_plan_launch *out, int32_t max_out, int32_t *n_out) {
    if ((!nodes && n > 0) || n < 0 || !n_out) return LK_EINVAL_ARG;
    Planner P;
    P.nodes = nodes; P.n = n;
    P.plan();
    const std::vector<lk_plan_launch> L = P.launches(chain_mode, 8);
    *n_out = (int32_t)L.size();
    if (out) {
        if ((int32_t)L.size() > max_out) return LK_EBOUNDS;
        for (size_t i = 0; i < L.size(); i++) out[i] = L[i];
    }
    return LK_OK;
}

static float bits_f(int32_t b) { float f; memcpy(&f, &b, 4); return f; }

// one node on its own: the `when (node.op)` of GGMLComputeOps.computeNode, core/GGMLComputeOps.kt:2515-2550
static lk_status run_node(lk_ctx *ctx, const lk_graph_node &nd) {
    switch (nd.op) {
        case LK_OP_NONE: return LK_OK;
        case LK_OP_MUL_MAT: return lk_compute_matmul(ctx, &nd.src[0], &nd.src[1], &nd.dst);                 // computeMulMat, :2647-2652
        case LK_OP_ADD: return lk_compute_add(ctx, &nd.src[0], &nd.src[1], &nd.dst);
        case LK_OP_MUL: return lk_compute_mul(ctx, &nd.src[0], &nd.src[1], &nd.dst);
        case LK_OP_SILU: return lk_compute_silu(ctx, &nd.src[0], &nd.dst);
        case LK_OP_RMS_NORM: return lk_compute_rms_norm(ctx, &nd.src[0], nd.n_src > 1 ? &nd.src[1] : nullptr, bits_f(nd.op_params[0]), &nd.dst);
        case LK_OP_SOFT_MAX: return lk_compute_softmax(ctx, &nd.src[0], &nd.dst);
        case LK_OP_SCALE: return lk_compute_scale(ctx, &nd.src[0], bits_f(nd.op_params[0]), &nd.dst);
        case LK_OP_ROPE: return lk_compute_rope(ctx, &nd.src[0], &nd.src[2], &nd.src[1], nd.op_params[0], &nd.dst);
        case LK_OP_CPY: return lk_compute_kv_append(ctx, &nd.src[0], &nd.src[1], &nd.src[2]);
        case LK_OP_FLASH_ATTN: return lk_compute_attention_decode(ctx, &nd.src[0], &nd.src[1], &nd.src[2], &nd.src[3], bits_f(nd.op_params[0]), &nd.dst);
        default: return lk::fail(ctx, LK_ENOTIMPL, "Operation not implemented on the CUDA backend");        // :2548
    }
}

static void fill_fusion(const lk_graph_node *nodes, const lk_plan_launch &L, lk_matmul_fusion *f, bool store_b) {
    memset(f, 0, sizeof(*f));
    const lk_graph_node &node = nodes[L.node];
    if (L.rms >= 0) {
        const lk_graph_node &r = nodes[L.rms];
        f->norm_x = &r.src[0];
        f->norm_w = r.n_src > 1 ? &r.src[1] : nullptr;
        f->eps = bits_f(r.op_params[0]);
        f->flags = store_b ? 0 : LK_FUSE_SKIP_B_STORE;
    }
    if (L.add >= 0) {
        const lk_graph_node &a = nodes[L.add];
        f->residual = key_of(a.src[0]) == key_of(node.dst) ? &a.src[1] : &a.src[0];
        f->sum_dst = &a.dst;
    }
    f->n_extra = L.n_extra;
    for (int q = 0; q < L.n_extra; q++) { f->a_extra[q] = &nodes[L.extra[q]].src[0]; f->dst_extra[q] = &nodes[L.extra[q]].dst; }
    if (L.mul >= 0) { f->silu_dst = &nodes[L.silu].dst; f->mul_dst = &nodes[L.mul].dst; }
}

// one planned launch on its own, with the unfused ops in graph order when the library declines the shape
static lk_status run_planned(lk_ctx *ctx, const lk_graph_node *nodes, const lk_plan_launch &L, std::set<int> &stored) {
    lk_status st;
    if (L.kind == LK_PLAN_NODE) return run_node(ctx, nodes[L.node]);
    if (L.kind == LK_PLAN_ATTN_STEP) {
        const lk_graph_node &rq = nodes[L.attn[0]], &rk = nodes[L.attn[1]], &ck = nodes[L.attn[2]], &cv = nodes[L.attn[3]], &fa = nodes[L.attn[4]];
        st = lk_compute_attention_step(ctx, &rq.src[0], &rk.src[0], &cv.src[0], &rq.src[2], &rq.src[1], rq.op_params[0], &ck.src[1], &cv.src[1],
                                       bits_f(fa.op_params[0]), &rq.dst, &rk.dst, &fa.dst);
        if (st != LK_ENOTIMPL) return st;
        for (int q = 0; q < 5; q++) if ((st = run_node(ctx, nodes[L.attn[q]])) != LK_OK) return st;
        return LK_OK;
    }
    const lk_graph_node &node = nodes[L.node];
    lk_matmul_fusion f;
    const bool store_b = L.rms >= 0 && !stored.count(L.rms);
    fill_fusion(nodes, L, &f, store_b);
    st = lk_compute_matmul_fused(ctx, &node.src[0], &node.src[1], &node.dst, &f);
    if (st == LK_OK) { if (L.rms >= 0) stored.insert(L.rms); return LK_OK; }
    if (st != LK_ENOTIMPL) return st;
    if (L.n_extra > 0) {
        // a declined group: every member (with its RMS_NORM prologue where it has one) on its own, then SILU / MUL
        std::vector<int> members{L.node};
        for (int q = 0; q < L.n_extra; q++) members.push_back(L.extra[q]);
        for (int m : members) {
            lk_plan_launch one;
            memset(&one, 0xFF, sizeof(one));
            one.kind = L.rms >= 0 ? LK_PLAN_FUSED : LK_PLAN_NODE; one.node = m; one.rms = L.rms; one.n_extra = 0;
            if ((st = run_planned(ctx, nodes, one, stored)) != LK_OK) return st;
        }
        if (L.mul >= 0) {
            if ((st = run_node(ctx, nodes[L.silu])) != LK_OK) return st;
            if ((st = run_node(ctx, nodes[L.mul])) != LK_OK) return st;
        }
        return LK_OK;
    }
    // a declined single launch: the unfused ops, in graph order
    if (L.rms >= 0 && !stored.count(L.rms)) {
        if ((st = run_node(ctx, nodes[L.rms])) != LK_OK) return st;
        stored.insert(L.rms);
    }
    if ((st = run_node(ctx, node)) != LK_OK) return st;
    if (L.add >= 0) return run_node(ctx, nodes[L.add]);
    return LK_OK;
}

// Host I/O of a graph: which launch each copy goes before (uploads: the first launch that reads the region) or after
// (downloads: the last launch that writes it).  Launches are numbered in issue order; a chained run counts as one.
namespace {
struct IoSchedule {
    std::vector<std::vector<int>> before, after;     // per launch slot: indices into io[]
    std::vector<int> at_end;                         // downloads of regions no node writes
};
bool node_reads(const lk_graph_node &nd, const Span &r) {
    for (int s = 0; s < nd.n_src && s < 4; s++) if (overlaps(span_of(nd.src[s]), r)) return true;
    return false;
}
bool node_writes(const lk_graph_node &nd, const Span &r) {
    if (nd.op == LK_OP_CPY) return nd.n_src > 1 && overlaps(span_of(nd.src[1]), r);   // the cache rows
    return nd.op != LK_OP_NONE && overlaps(span_of(nd.dst), r);
}
IoSchedule schedule_io(const lk_graph_node *nodes, const std::vector<std::vector<int>> &covered, const lk_host_io *io, int n_io) {
    IoSchedule S;
    const int nl = (int)covered.size();
    S.before.resize(nl + 1);
    S.after.resize(nl + 1);
    for (int k = 0; k < n_io; k++) {
        const Span r{(const void *)io[k].buf, io[k].offset, io[k].offset + io[k].bytes};
        int at = -1;
        for (int l = 0; l < nl; l++) {
            bool hit = false;
            for (int j : covered[l]) hit = hit || (io[k].direction == LK_IO_UPLOAD ? node_reads(nodes[j], r) : node_writes(nodes[j], r));
            if (!hit) continue;
            at = l;
            if (io[k].direction == LK_IO_UPLOAD) break;     // first reader; downloads keep the last writer
        }
        if (io[k].direction == LK_IO_UPLOAD) S.before[at < 0 ? 0 : at].push_back(k);
        else if (at < 0) S.at_end.push_back(k);
        else S.after[at].push_back(k);
    }
    return S;
}
lk_status issue_io(const lk_host_io &c) {
    return c.direction == LK_IO_UPLOAD ? lk_buf_upload_async(c.buf, c.offset, c.host, c.bytes)
                                       : lk_buf_download_async(c.buf, c.offset, c.host, c.bytes);
}
std::vector<int> launch_nodes(const lk_plan_launch &L) {
    std::vector<int> v;
    if (L.kind == LK_PLAN_ATTN_STEP) { for (int q = 0; q < 5; q++) v.push_back(L.attn[q]); return v; }
    v.push_back(L.node);
    if (L.kind == LK_PLAN_FUSED) {
        if (L.rms >= 0) v.push_back(L.rms);
        if (L.add >= 0) v.push_back(L.add);
        for (int q = 0; q < L.n_extra; q++) v.push_back(L.extra[q]);
        if (L.mul >= 0) { v.push_back(L.silu); v.push_back(L.mul); }
    }
    return v;
}
}  // namespace

lk_status lk_graph_compute_io(lk_ctx *ctx, const lk_graph_node *nodes, int32_t n, int32_t flags, const lk_host_io *io, int32_t n_io) {
    if (!ctx || (!nodes && n > 0) || n < 0 || n_io < 0 || (!io && n_io > 0)) return lk::fail(ctx, LK_EINVAL_ARG, "lk_graph_compute: bad argument");
    for (int k = 0; k < n_io; k++)
        if (!io[k].buf || !io[k].host || (io[k].direction != LK_IO_UPLOAD && io[k].direction != LK_IO_DOWNLOAD))
            return lk::fail(ctx, LK_EINVAL_ARG, "lk_graph_compute_io: a copy needs a device buffer, a host address and a direction");
    lk_status st;
    if (flags & LK_GRAPH_NO_FUSION) {
        std::vector<std::vector<int>> covered(n);
        for (int i = 0; i < n; i++) covered[i].push_back(i);
        const IoSchedule S = schedule_io(nodes, covered, io, n_io);
        for (int i = 0; i < n; i++) {
            for (int k : S.before[i]) if ((st = issue_io(io[k])) != LK_OK) return st;
            if ((st = run_node(ctx, nodes[i])) != LK_OK) return st;
            for (int k : S.after[i]) if ((st = issue_io(io[k])) != LK_OK) return st;
        }
        for (int k : S.before[n]) if ((st = issue_io(io[k])) != LK_OK) return st;
        for (int k : S.at_end) if ((st = issue_io(io[k])) != LK_OK) return st;
        return LK_OK;
    }
    Planner P;
    P.nodes = nodes; P.n = n;
    P.plan();
    const int chain_mode = (flags & LK_GRAPH_CHAIN_ALWAYS) ? 1 : (flags & LK_GRAPH_CHAIN_NEVER) ? 0 : -1;
    const std::vector<lk_plan_launch> L = P.launches(chain_mode, 8);
    // launch slots: one per unchained launch, one per chained run
    std::vector<std::pair<size_t, size_t>> slots;
    for (size_t i = 0; i < L.size();) {
        size_t e = i + 1;
        if (L[i].chain >= 0) while (e < L.size() && L[e].chain == L[i].chain) e++;
        slots.push_back({i, e});
        i = e;
    }
    std::vector<std::vector<int>> covered(slots.size());
    for (size_t q = 0; q < slots.size(); q++)
        for (size_t i = slots[q].first; i < slots[q].second; i++)
            for (int j : launch_nodes(L[i])) covered[q].push_back(j);
    const IoSchedule S = schedule_io(nodes, covered, io, n_io);
    std::set<int> stored;
    for (size_t q = 0; q < slots.size(); q++) {
        const size_t i = slots[q].first, e = slots[q].second;
        for (int k : S.before[q]) if ((st = issue_io(io[k])) != LK_OK) return st;
        if (L[i].chain < 0) {
            if ((st = run_planned(ctx, nodes, L[i], stored)) != LK_OK) return st;
        } else {
            // consecutive dependent launches as one lk_compute_matmul_chain; one by one when the library declines
            const std::set<int> saved = stored;
            std::vector<lk_matmul_fusion> fus(e - i);
            std::vector<lk_matmul_link> links(e - i);
            for (size_t z = i; z < e; z++) {
                const bool store_b = L[z].rms >= 0 && !stored.count(L[z].rms);
                const bool has = L[z].kind == LK_PLAN_FUSED;
                if (has) fill_fusion(nodes, L[z], &fus[z - i], store_b);
                if (L[z].rms >= 0) stored.insert(L[z].rms);
                links[z - i].a = &nodes[L[z].node].src[0];
                links[z - i].b = &nodes[L[z].node].src[1];
                links[z - i].dst = &nodes[L[z].node].dst;
                links[z - i].fusion = has ? &fus[z - i] : nullptr;
            }
            st = lk_compute_matmul_chain(ctx, links.data(), (int32_t)links.size());
            if (st == LK_ENOTIMPL) {
                stored = saved;
                for (size_t z = i; z < e; z++) if ((st = run_planned(ctx, nodes, L[z], stored)) != LK_OK) return st;
            } else if (st != LK_OK) {
                return st;
            }
        }
        for (int k : S.after[q]) if ((st = issue_io(io[k])) != LK_OK) return st;
    }
    for (int k : S.before[slots.size()]) if ((st = issue_io(io[k])) != LK_OK) return st;
    for (int k : S.at_end) if ((st = issue_io(io[k])) != LK_OK) return st;
    return LK_OK;
}

lk_status lk_graph_compute(lk_ctx *ctx, const lk_graph_node *nodes, int32_t n, int32_t flags) {
    return lk_graph_compute_io(ctx, nodes, n, flags, nullptr, 0);
}

}  // extern "C"
