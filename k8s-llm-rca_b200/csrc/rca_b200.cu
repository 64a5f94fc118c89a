// rca_b200.cu - C ABI (include/rca_b200.h) over the kernels in rca_kernels.cuh.
//
// Host side of the engine: uploads a snapshot to HBM, turns a plan into one launch per hop, keeps
// the scratch arenas (frontier entries, output rows, tile states) and hands results back.  There
// is no CPU path: every entry point that computes runs kernels on the device it was opened on.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "rca_b200.h"
#include "rca_kernels.cuh"

using namespace rca;

static thread_local std::string g_open_error;

// ------------------------------------------------------------------------------------------------
// One plan, prepared on the host: its table (kind boundaries, one kind-action row per level,
// node-block -> kind table, one arrival row per level) and, per level, how to launch it.
struct Sub { int t, d, fl, aux; };    // fl / aux: what LevelParams::unit_fl / unit_aux say about the step's far kind
struct PreparedPlan {
    const rca_plan_t *pl = nullptr;
    bool varlen = false, fused = false, arr_prune = false;
    bool walk = false;                           // typed chain with little fan-out: one chain_walk launch
    long long fan = 0;                           // ... and the most rows one seed can have
    int L = 0, mode = kModeChain;
    std::vector<std::vector<Sub>> subs;          // typed sub-launches per level (split path)
    std::vector<char> level_typed, level_dead;
    std::vector<std::vector<char>> cont_kind;    // [level][kind]: a path may CONTINUE through a node of this kind at this level
    int wide_level = -1;                         // level whose frontier is stored as wide entries (rca_kernels.cuh), or -1
};

struct PlanResult {
    bool valid = false;
    int levels = 0;
    int row_len[RCA_MAX_HOPS + 1] = {0};
    long long rows[RCA_MAX_HOPS + 1] = {0};
    unsigned long long base[RCA_MAX_HOPS + 1] = {0};
};

struct rca_engine {
    int device = 0;
    int sm_count = 148;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ev_lv[RCA_MAX_HOPS + 2] = {nullptr};
    std::string err;
    // graph
    int n_nodes = 0, n_kinds = 0, n_etypes = 0;
    long long n_rels = 0, n_und = 0, n_states = 0;
    std::vector<int> kind_start, et_src, et_dst, et_cls;
    std::vector<unsigned char> kind_lut;   // node id block -> kind of the block's first id
    int lut_shift = 0;
    std::vector<long long> typed_out_base, typed_in_base;
    int2 *out_rng = nullptr, *in_rng = nullptr, *und_rng = nullptr, *in_adj = nullptr, *und_adj = nullptr;
    int2 *typed_out = nullptr, *typed_in = nullptr;
    int *out_dst = nullptr, *st_off = nullptr;
    StRec *st_rec = nullptr;              // HasState intervals, one 32 B record each
    StHead *st_head = nullptr;            // first interval + count + offset per node that can have one (rca_kernels.cuh)
    StMore *st_more = nullptr;            // second and third interval of the nodes that have them
    int n_head = 0;
    unsigned char *rel_et = nullptr;      // edge type per relationship = per out_dst slot (only when there are <= 256 edge types)
    unsigned char *in_et = nullptr, *und_et = nullptr;   // ... and per in_adj / und_adj slot
    std::vector<int> et_max_out, et_max_in;   // largest number of relationships of the type at one start / end node
    // edge types no node has two of (on that side): per node of the near kind, its (neighbour, relationship) or
    // (-1, -1), instead of a range into an adjacency array; base -1 = the type is not like that
    int2 *unit_next = nullptr;
    size_t graph_bytes = 0;               // device memory held by the snapshot (adjacency, typed tables, intervals, the tables above)
    std::vector<long long> unit_out_base, unit_in_base;
    // scratch
    int4 *frontier = nullptr;
    unsigned long long fcap = 0;
    int *out = nullptr;
    unsigned long long ocap = 0;
    // double buffering of the output arena (rca_set_double_buffer): every submission takes the other arena, so
    // the rows of the previous one can still be travelling to the host (rca_result_copy_async) while it runs
    bool double_buffer = false;
    int *out_alt = nullptr;
    unsigned long long ocap_alt = 0;
    int arena = 0;                        // which of the two is `out`
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copied[2] = {nullptr, nullptr};   // last asynchronous copy out of arena 0 / 1
    bool copies_pending[2] = {false, false};
    ulonglong2 *tstate = nullptr;         // fused path: tile-state words, tagged with `epoch`; zeroed when (re)allocated
    unsigned long long tcap = 0;
    unsigned epoch = 0;
    uint2 *tcount = nullptr;              // split path: per-tile counts
    unsigned long long ccap = 0;
    ulonglong2 *gsum = nullptr, *gscan = nullptr;   // split path: group sums (zeroed per call) and their scan
    unsigned long long gcap = 0;
    long long fused_max_seeds = 16384;    // up to this many seeds a variable-length plan runs one fused launch per level
    long long chain_walk_fanout = 64;     // typed chains whose per-seed fan-out is at most this run as one depth-first launch
    bool unit_steps = true;               // typed steps with at most one candidate per path use count_unit / fill_unit
    bool wide_entries = false;            // the level before a final unit level stores whole paths (rca_set_wide_entries);
                                          // measured on config 3: level 2 0.095 -> 0.102 ms, level 3 0.189 -> 0.187 ms - off
    long long chain_fused_max_seeds = 1 << 20;   // typed chains have uniform tiles: the fused launch suits them far longer
    Ctl *ctl = nullptr;
    Ctl *h_ctl = nullptr;                 // pinned
    unsigned char *ptab = nullptr, *h_ptab = nullptr;   // plan table (rca_kernels.cuh: kPt*)
    int *seeds_buf = nullptr;
    long long seeds_cap = 0;
    long long *ts_buf = nullptr;
    long long ts_cap = 0;
    unsigned char *msg_bytes = nullptr;   // message arena for rca_text_contains
    long long *msg_off = nullptr;
    std::vector<int> msg_node;            // message index -> node id (host)
    long long n_msgs = 0;
    void *io_buf = nullptr;               // staging for the state and text queries with host buffers
    size_t io_cap = 0;
    // rca_expand with few host seeds (the per-incident statements): plan table, control block, timestamps and seeds lie side by
    // side in one device block with a pinned mirror, so that a call uploads ONCE
    unsigned char *small_dev = nullptr, *small_host = nullptr;
    void *io_host = nullptr;              // ... and its pinned mirror: a call's inputs go up in ONE copy and its outputs come down
    size_t io_host_cap = 0;               // in ONE (every small copy is a DMA of its own, 5-8 us each, and they queue up)
    // rca_expand_many: per-plan tables and control blocks
    unsigned char *m_ptab = nullptr, *mh_ptab = nullptr;
    Ctl *m_ctl = nullptr, *mh_ctl = nullptr;
    size_t many_cap = 0;
    std::vector<cudaEvent_t> ev_many;
    struct Occ { const void *kernel; size_t dyn; int grid; };
    std::vector<Occ> occ_cache;           // launch grid per (kernel, dynamic shared memory)
    struct CachedPlan {
        rca_plan_t plan;
        bool fused;
        PreparedPlan pp;
        std::vector<unsigned char> table;
    };
    std::vector<CachedPlan> plan_cache;   // prepared plans of earlier calls (prepare_plan): the schema analysis is done once
    size_t plan_cache_next = 0;
    // rca_expand_many: plans whose output is bounded on the host run beside the others on these streams
    std::vector<cudaStream_t> side;
    std::vector<cudaEvent_t> ev_side;
    cudaEvent_t ev_fork = nullptr;
    // count exchange over peer memory (rca_peer_*)
    int peer_rank = -1, peer_world = 0;
    PeerSlot *peer_own = nullptr;         // [2][world] this rank's slots, written by the peers
    PeerSlot *peer_slots[8] = {nullptr};  // every rank's slot array (own one included), as mapped into this process
    long long *peer_out = nullptr, *peer_out_h = nullptr;   // [world] device / pinned
    unsigned *peer_err = nullptr;
    unsigned long long peer_seq = 0;
    bool peer_collected = true;           // peer_out holds the values of exchange peer_seq
    unsigned long long peer_timeout_ns = 2000000000ull;
    // row gather over peer memory (rca_gather_*): the window rows are pushed into
    int *gather_own = nullptr;            // this rank's window (only a rank that is a destination has one)
    unsigned long long gather_own_words = 0;
    int *gather_dst = nullptr;            // the destination rank's window as mapped into this process
    unsigned long long gather_dst_words = 0;
    int gather_dst_rank = -1;
    bool gather_dst_mapped = false;       // gather_dst came from cudaIpcOpenMemHandle (not this rank's own window)
    cudaEvent_t ev_g0 = nullptr, ev_g1 = nullptr;
    // rca_expand_many: a submission seen twice with the same plans, seed buffers and arenas is replayed as a CUDA graph
    // (one cudaGraphLaunch instead of ~25 stream operations: the host side of a 0.5 ms step is 0.2 ms of enqueueing)
    struct ManyGraph {
        std::vector<unsigned char> key;
        cudaGraphExec_t exec = nullptr;
        std::vector<int> launches, leader;
        unsigned long long stamp = 0;
    };
    std::vector<ManyGraph> many_graphs;
    std::vector<std::vector<unsigned char>> many_seen;   // keys of the last few eager submissions: one seen again is captured
    unsigned long long many_stamp = 0;
    bool use_graphs = false;              // measured on config 3: a replay is 4 % (one GPU) to 7 % (eight processes on one
                                          // host) SLOWER than the eager enqueue - the step is bound by its kernels' dependent
                                          // chain, not by the host, and the graph's event / memcpy nodes cost more than they save
    bool capturing = false;               // stream operations are being captured: timing events are recorded as external
    // results of the last submission (one per plan) and the one the accessors address
    std::vector<PlanResult> results;
    int selected = 0;
};

static void drop_graphs(rca_engine *e);

static int fail(rca_engine *e, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (e) e->err = buf; else g_open_error = buf;
    return code;
}

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess)                                                                     \
            return fail(e, _e == cudaErrorMemoryAllocation ? RCA_ERR_NOMEM : RCA_ERR_CUDA,          \
                        "%s failed: %s", #call, cudaGetErrorString(_e));                            \
    } while (0)

template <typename T>
static cudaError_t upload(T **dst, const void *src, size_t count) {
    *dst = nullptr;
    size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
    cudaError_t r = cudaMalloc((void **)dst, bytes);
    if (r != cudaSuccess) return r;
    if (count) r = cudaMemcpy(*dst, src, count * sizeof(T), cudaMemcpyHostToDevice);
    return r;
}

extern "C" const char *rca_version(void) { return "rca_b200 0.1 sm_100a"; }

extern "C" const char *rca_last_error(const rca_engine_t *e) { return e ? e->err.c_str() : g_open_error.c_str(); }

extern "C" int rca_open(const rca_snapshot_t *s, int device, rca_engine_t **out) {
    rca_engine *e = nullptr;
    if (!s || !out) return fail(e, RCA_ERR_ARG, "rca_open: null argument");
    *out = nullptr;
    if (s->n_kinds < 1 || s->n_kinds > RCA_MAX_KINDS) return fail(e, RCA_ERR_ARG, "rca_open: n_kinds %d not in 1..%d", s->n_kinds, RCA_MAX_KINDS);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(e, RCA_ERR_NODEVICE, "rca_open: no CUDA device is visible; this library has no CPU path");
    if (device < 0 || device >= ndev) return fail(e, RCA_ERR_ARG, "rca_open: device %d not in 0..%d", device, ndev - 1);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return fail(e, RCA_ERR_CUDA, "rca_open: cannot query device");
    if (prop.major != 10)
        return fail(e, RCA_ERR_NODEVICE, "rca_open: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    std::unique_ptr<rca_engine> up(new rca_engine);
    rca_engine *g = up.get();
#define OCK(call)                                                                                  \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            int rc = fail(nullptr, _e == cudaErrorMemoryAllocation ? RCA_ERR_NOMEM : RCA_ERR_CUDA,  \
                          "rca_open: %s failed: %s", #call, cudaGetErrorString(_e));                \
            rca_close(up.release());                                                               \
            return rc;                                                                             \
        }                                                                                          \
    } while (0)
    g->device = device;
    g->sm_count = prop.multiProcessorCount;
    if (const char *v = std::getenv("RCA_B200_GRAPHS")) g->use_graphs = v[0] != '0';       // measurements: replay on / off
    OCK(cudaSetDevice(device));
    {
        int least = 0, greatest = 0;
        OCK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        OCK(cudaStreamCreateWithPriority(&g->own_stream, cudaStreamNonBlocking, greatest));
    }
    g->stream = g->own_stream;
    OCK(cudaEventCreate(&g->ev0));
    OCK(cudaEventCreate(&g->ev1));
    for (auto &ev : g->ev_lv) OCK(cudaEventCreate(&ev));
    g->n_nodes = s->n_nodes; g->n_kinds = s->n_kinds; g->n_etypes = s->n_etypes;
    g->n_rels = s->n_rels; g->n_und = s->n_und; g->n_states = s->n_states;
    g->kind_start.assign(s->kind_start, s->kind_start + s->n_kinds + 1);
    while (((long long)std::max(s->n_nodes, 1) - 1) >> g->lut_shift >= kLutEntries) g->lut_shift++;
    g->kind_lut.assign(kLutEntries, 0);
    for (int b = 0, k = 0; b < kLutEntries; b++) {
        const long long v = (long long)b << g->lut_shift;
        while (k + 1 < s->n_kinds && s->kind_start[k + 1] <= v) k++;
        g->kind_lut[b] = (unsigned char)k;
    }
    g->et_src.assign(s->et_src, s->et_src + s->n_etypes);
    g->et_dst.assign(s->et_dst, s->et_dst + s->n_etypes);
    g->et_cls.assign(s->et_cls, s->et_cls + s->n_etypes);
    g->typed_out_base.assign(s->typed_out_base, s->typed_out_base + s->n_etypes + 1);
    g->typed_in_base.assign(s->typed_in_base, s->typed_in_base + s->n_etypes + 1);
    const size_t V1 = (size_t)s->n_nodes + 1;
    OCK(upload(&g->out_rng, s->out_rng, V1));
    OCK(upload(&g->in_rng, s->in_rng, V1));
    OCK(upload(&g->und_rng, s->und_rng, V1));
    OCK(upload(&g->out_dst, s->out_dst, (size_t)s->n_rels));
    OCK(upload(&g->in_adj, s->in_adj, (size_t)s->n_rels));
    OCK(upload(&g->und_adj, s->und_adj, (size_t)s->n_und));
    OCK(upload(&g->typed_out, s->typed_out, (size_t)s->n_typed_out));
    OCK(upload(&g->typed_in, s->typed_in, (size_t)s->n_typed_in));
    OCK(upload(&g->st_off, s->st_off, V1));
    {
        std::vector<StRec> rec((size_t)s->n_states);
        for (size_t j = 0; j < rec.size(); j++) rec[j] = StRec{s->st_tmin[j], s->st_tmax[j], s->st_node[j], {0, 0, 0}};
        OCK(upload(&g->st_rec, rec.data(), rec.size()));
        int last = -1;                                 // entity kinds come first: the table stops at the last node with an interval
        for (int v = 0; v < s->n_nodes; v++) if (s->st_off[v + 1] > s->st_off[v]) last = v;
        g->n_head = last + 1;
        std::vector<StHead> head((size_t)g->n_head);
        auto t44 = [](long long t) { return (unsigned long long)std::min<long long>(std::max<long long>(t, 0), (long long)kT44); };
        for (int v = 0; v < g->n_head; v++) {
            const int b = s->st_off[v], c = s->st_off[v + 1] - b;
            if (!c) { head[v] = StHead{1ull, 0ull}; continue; }
            const unsigned long long node = (unsigned)s->st_node[b], more = (unsigned long long)std::min(c - 1, 255);
            head[v] = StHead{t44(s->st_tmin[b]) | (node & 0xfffffull) << 44, t44(s->st_tmax[b]) | (node >> 20) << 44 | more << 56};
        }
        OCK(upload(&g->st_head, head.data(), head.size()));
        std::vector<StMore> more((size_t)g->n_head, StMore{1ull, 0ull, 1ull, 0ull});
        for (int v = 0; v < g->n_head; v++) {
            const int b = s->st_off[v], c = s->st_off[v + 1] - b;
            auto pack = [&](int j, unsigned long long &a_, unsigned long long &b_) {
                const unsigned long long node = (unsigned)s->st_node[j];
                a_ = t44(s->st_tmin[j]) | (node & 0xfffffull) << 44;
                b_ = t44(s->st_tmax[j]) | (node >> 20) << 44;
            };
            if (c >= 2) pack(b + 1, more[v].a1, more[v].b1);
            if (c >= 3) pack(b + 2, more[v].a2, more[v].b2);
        }
        OCK(upload(&g->st_more, more.data(), more.size()));
    }
    if (s->n_etypes <= 256 && s->rel_et) {
        std::vector<unsigned char> et((size_t)s->n_rels);
        for (size_t j = 0; j < et.size(); j++) et[j] = (unsigned char)s->rel_et[j];
        OCK(upload(&g->rel_et, et.data(), et.size()));
        std::vector<unsigned char> ie((size_t)s->n_rels), ue((size_t)s->n_und);
        for (size_t j = 0; j < ie.size(); j++) ie[j] = et[(size_t)(s->in_adj[2 * j + 1] & kRelMask)];
        for (size_t j = 0; j < ue.size(); j++) ue[j] = et[(size_t)(s->und_adj[2 * j + 1] & kRelMask)];
        OCK(upload(&g->in_et, ie.data(), ie.size()));
        OCK(upload(&g->und_et, ue.data(), ue.size()));
    }
    g->et_max_out.assign(s->n_etypes, 0);
    g->et_max_in.assign(s->n_etypes, 0);
    for (int t = 0; t < s->n_etypes; t++) {
        for (long long j = s->typed_out_base[t]; j < s->typed_out_base[t + 1]; j++)
            g->et_max_out[t] = std::max(g->et_max_out[t], s->typed_out[2 * j + 1] - s->typed_out[2 * j]);
        for (long long j = s->typed_in_base[t]; j < s->typed_in_base[t + 1]; j++)
            g->et_max_in[t] = std::max(g->et_max_in[t], s->typed_in[2 * j + 1] - s->typed_in[2 * j]);
    }
    {
        std::vector<int2> nx;
        g->unit_out_base.assign(s->n_etypes, -1);
        g->unit_in_base.assign(s->n_etypes, -1);
        for (int t = 0; t < s->n_etypes; t++) {
            if (g->et_max_out[t] <= 1) {
                g->unit_out_base[t] = (long long)nx.size();
                for (long long j = s->typed_out_base[t]; j < s->typed_out_base[t + 1]; j++) {
                    const int b = s->typed_out[2 * j];
                    nx.push_back(s->typed_out[2 * j + 1] > b ? make_int2(s->out_dst[b], b) : make_int2(-1, -1));
                }
            }
            if (g->et_max_in[t] <= 1) {
                g->unit_in_base[t] = (long long)nx.size();
                for (long long j = s->typed_in_base[t]; j < s->typed_in_base[t + 1]; j++) {
                    const int b = s->typed_in[2 * j];
                    nx.push_back(s->typed_in[2 * j + 1] > b ? make_int2(s->in_adj[2 * (size_t)b], s->in_adj[2 * (size_t)b + 1] & kRelMask)
                                                            : make_int2(-1, -1));
                }
            }
        }
        OCK(upload(&g->unit_next, nx.data(), nx.size()));
        g->graph_bytes = 3 * V1 * sizeof(int2) + (size_t)s->n_rels * (sizeof(int) + sizeof(int2)) + (size_t)s->n_und * sizeof(int2) +
                         ((size_t)s->n_typed_out + (size_t)s->n_typed_in) * sizeof(int2) + V1 * sizeof(int) +
                         (size_t)s->n_states * sizeof(StRec) + (g->rel_et ? 2 * (size_t)s->n_rels + (size_t)s->n_und : 0) +
                         (size_t)g->n_head * (sizeof(StHead) + sizeof(StMore)) + nx.size() * sizeof(int2);
    }
    OCK(cudaMalloc((void **)&g->ctl, sizeof(Ctl)));
    OCK(cudaMallocHost((void **)&g->h_ctl, sizeof(Ctl)));
    OCK(cudaMalloc((void **)&g->ptab, kPtBytes));
    OCK(cudaMallocHost((void **)&g->h_ptab, kPtBytes));
#undef OCK
    *out = up.release();
    return RCA_OK;
}

extern "C" int rca_close(rca_engine_t *e) {
    if (!e) return RCA_OK;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    for (auto &g : e->many_graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
    void *dev[] = {e->out_rng, e->in_rng, e->und_rng, e->in_adj, e->und_adj, e->typed_out, e->typed_in, e->out_dst,
                   e->st_off, e->st_rec, e->st_head, e->st_more, e->unit_next, e->rel_et, e->in_et, e->und_et, e->frontier, e->out, e->tstate, e->tcount, e->gsum, e->gscan, e->ctl, e->ptab,
                   e->seeds_buf, e->ts_buf, e->io_buf, e->msg_bytes, e->msg_off, e->m_ptab, e->m_ctl};
    for (void *p : dev) if (p) cudaFree(p);
    if (e->io_host) cudaFreeHost(e->io_host);
    if (e->small_host) cudaFreeHost(e->small_host);
    if (e->small_dev) cudaFree(e->small_dev);
    if (e->h_ctl) cudaFreeHost(e->h_ctl);
    if (e->h_ptab) cudaFreeHost(e->h_ptab);
    if (e->mh_ptab) cudaFreeHost(e->mh_ptab);
    if (e->mh_ctl) cudaFreeHost(e->mh_ctl);
    if (e->copy_stream) { cudaStreamSynchronize(e->copy_stream); cudaStreamDestroy(e->copy_stream); }
    for (auto ev : e->ev_copied) if (ev) cudaEventDestroy(ev);
    if (e->out_alt) cudaFree(e->out_alt);
    for (int r = 0; r < e->peer_world; r++)
        if (r != e->peer_rank && e->peer_slots[r]) cudaIpcCloseMemHandle(e->peer_slots[r]);
    if (e->peer_own) cudaFree(e->peer_own);
    if (e->peer_out) cudaFree(e->peer_out);
    if (e->peer_err) cudaFree(e->peer_err);
    if (e->peer_out_h) cudaFreeHost(e->peer_out_h);
    if (e->gather_dst_mapped && e->gather_dst) cudaIpcCloseMemHandle(e->gather_dst);
    if (e->gather_own) cudaFree(e->gather_own);
    if (e->ev_g0) cudaEventDestroy(e->ev_g0);
    if (e->ev_g1) cudaEventDestroy(e->ev_g1);
    for (auto ev : e->ev_many) cudaEventDestroy(ev);
    for (auto ev : e->ev_side) cudaEventDestroy(ev);
    for (auto st : e->side) cudaStreamDestroy(st);
    if (e->ev_fork) cudaEventDestroy(e->ev_fork);
    if (e->ev0) cudaEventDestroy(e->ev0);
    if (e->ev1) cudaEventDestroy(e->ev1);
    for (auto ev : e->ev_lv) if (ev) cudaEventDestroy(ev);
    if (e->own_stream) cudaStreamDestroy(e->own_stream);
    delete e;
    return RCA_OK;
}

extern "C" int rca_set_graphs(rca_engine_t *e, int on) {
    if (!e) return RCA_ERR_ARG;
    e->use_graphs = on != 0;
    if (!on) { cudaSetDevice(e->device); drop_graphs(e); }
    return RCA_OK;
}

extern "C" int rca_set_stream(rca_engine_t *e, uint64_t cuda_stream) {
    if (!e) return RCA_ERR_ARG;
    e->stream = cuda_stream ? (cudaStream_t)(uintptr_t)cuda_stream : e->own_stream;
    return RCA_OK;
}

template <typename T>
static int grow(rca_engine *e, T **buf, unsigned long long *cap, unsigned long long want) {
    if (want <= *cap && *buf) return RCA_OK;
    CK(cudaStreamSynchronize(e->stream));
    if (*buf) CK(cudaFree(*buf));
    *buf = nullptr;
    *cap = 0;
    CK(cudaMalloc((void **)buf, std::max<unsigned long long>(want, 1) * sizeof(T)));
    *cap = want;
    return RCA_OK;
}

extern "C" int64_t rca_graph_bytes(const rca_engine_t *e) { return e ? (int64_t)e->graph_bytes : 0; }

extern "C" int rca_set_unit_steps(rca_engine_t *e, int on) {
    if (!e) return RCA_ERR_ARG;
    e->unit_steps = on != 0;
    e->plan_cache.clear();
    drop_graphs(e);
    return RCA_OK;
}

extern "C" int rca_set_wide_entries(rca_engine_t *e, int on) {
    if (!e) return RCA_ERR_ARG;
    e->wide_entries = on != 0;
    e->plan_cache.clear();
    drop_graphs(e);
    return RCA_OK;
}

extern "C" int rca_reserve(rca_engine_t *e, int64_t frontier_entries, int64_t output_words) {
    if (!e || frontier_entries < 0 || output_words < 0) return RCA_ERR_ARG;
    CK(cudaSetDevice(e->device));
    int rc = grow(e, &e->frontier, &e->fcap, (unsigned long long)frontier_entries);
    if (rc) return rc;
    e->results.clear();
    return grow(e, &e->out, &e->ocap, (unsigned long long)output_words);
}

// schema-level reachability: bit (r-1) of reach[k] says a walk of 1..r hops in the plan's
// direction, through kinds the plan does not exclude, leads from kind k to dest_kind
static void reach_bits(const rca_engine *e, const rca_plan_t *pl, unsigned *reach) {
    const int nk = e->n_kinds;
    bool constrained = false;                       // a required kind at position 0 only restricts the seeds
    for (int i = 1; i <= pl->max_hops; i++) constrained |= pl->pos_kind[i] >= 0;
    if (!pl->prune || pl->dest_kind < 0 || constrained) {
        for (int k = 0; k < nk; k++) reach[k] = 0xffu;
        return;
    }
    std::vector<unsigned char> in_d(nk, 0), nxt(nk, 0);
    for (int k = 0; k < nk; k++) reach[k] = 0;
    if (pl->kind_excl[pl->dest_kind]) return;
    auto edge = [&](int t, int &a, int &b, int flip) {   // a -> b in walk direction
        a = flip ? e->et_dst[t] : e->et_src[t];
        b = flip ? e->et_src[t] : e->et_dst[t];
    };
    for (int r = 1; r <= pl->max_hops; r++) {
        std::fill(nxt.begin(), nxt.end(), 0);
        for (int t = 0; t < e->n_etypes; t++) {
            if (pl->entity_only && e->et_cls[t]) continue;
            for (int flip = 0; flip < 2; flip++) {
                if (pl->direction == RCA_DIR_OUT && flip) continue;
                if (pl->direction == RCA_DIR_IN && !flip) continue;
                int a, b;
                edge(t, a, b, flip);
                bool hit = (b == pl->dest_kind) || (r > 1 && in_d[b] && !pl->kind_excl[b]);
                if (hit) nxt[a] = 1;
            }
        }
        for (int k = 0; k < nk; k++) {
            in_d[k] = in_d[k] | nxt[k];
            if (in_d[k]) reach[k] |= 1u << (r - 1);
        }
    }
}

// Arrival model shared by the pruning tables.  An arrival a = 2*t + d means the path's last hop
// followed a relationship of edge type t along (d = 0) or against (d = 1) its stored direction.
struct ArrivalModel {
    const rca_engine *e;
    const rca_plan_t *pl;
    int head(int a) const { return (a & 1) ? e->et_src[a >> 1] : e->et_dst[a >> 1]; }
    int tail(int a) const { return (a & 1) ? e->et_dst[a >> 1] : e->et_src[a >> 1]; }
    bool allowed(int a) const {
        const int t = a >> 1, d = a & 1;
        if (pl->entity_only && e->et_cls[t]) return false;
        if (pl->direction == RCA_DIR_OUT && d) return false;
        if (pl->direction == RCA_DIR_IN && !d) return false;
        return true;
    }
    // may arrival b follow arrival a ?  The one thing this knows beyond kinds: the next hop cannot go
    // straight back over the same edge type when that type is functional on that side (no node has
    // two such relationships) - the only candidate is the relationship just used, and a MATCH
    // pattern never reuses one.
    bool step_ok(int a, int b) const {
        if (!allowed(b) || tail(b) != head(a)) return false;
        if ((b >> 1) == (a >> 1) && (b & 1) != (a & 1)) {
            const int t = a >> 1;
            const int deg = (a & 1) ? e->et_max_out[t] : e->et_max_in[t];
            if (deg <= 1) return false;
        }
        return true;
    }
};

// rows[s][a] = 1 if a path that arrived by a and is now s hops long may still end at dest_kind within
// L hops (arrival-aware refinement of reach_bits).  Returns false (rows all 1) when it does not apply.
static bool arrival_reach(const rca_engine *e, const rca_plan_t *pl, int L, unsigned char *rows) {
    const int nt = e->n_etypes, A = 2 * nt;
    std::memset(rows, 1, (size_t)(RCA_MAX_HOPS + 1) * kArrBytes);
    bool constrained = false;
    for (int i = 1; i <= L; i++) constrained |= pl->pos_kind[i] >= 0;
    if (!pl->prune || pl->dest_kind < 0 || constrained || nt > 256 || !e->rel_et) return false;
    const ArrivalModel m{e, pl};
    std::vector<unsigned char> prev(A, 0), cur(A, 0);
    const bool dest_ok = !pl->kind_excl[pl->dest_kind];
    for (int r = 1; r <= L; r++) {
        for (int a = 0; a < A; a++) {
            unsigned char hit = prev[a];
            for (int b = 0; b < A && !hit; b++) {
                if (!m.step_ok(a, b)) continue;
                const int hb = m.head(b);
                if (hb == pl->dest_kind && dest_ok) hit = 1;
                else if (r > 1 && !pl->kind_excl[hb] && prev[b]) hit = 1;
            }
            cur[a] = hit;
        }
        unsigned char *row = rows + (size_t)(L - r) * kArrBytes;      // a path L - r hops long has r hops left
        for (int a = 0; a < A; a++) row[a] = cur[a];
        prev = cur;
    }
    return true;
}

static int validate_plan(rca_engine *e, const rca_plan_t *pl) {
    if (pl->with_states < 0 || pl->with_states > RCA_STATES_MASK) return fail(e, RCA_ERR_ARG, "plan: with_states %d not in 0..2", pl->with_states);
    if (pl->kind == RCA_PLAN_CHAIN) {
        if (pl->n_steps < 1 || pl->n_steps > RCA_MAX_HOPS) return fail(e, RCA_ERR_ARG, "chain plan: n_steps %d not in 1..%d", pl->n_steps, RCA_MAX_HOPS);
        int cur = -1;
        for (int i = 0; i < pl->n_steps; i++) {
            int t = pl->step_etype[i], d = pl->step_dir[i];
            if (t < 0 || t >= e->n_etypes) return fail(e, RCA_ERR_ARG, "chain plan: step %d edge type %d unknown", i, t);
            if (d != RCA_DIR_OUT && d != RCA_DIR_IN) return fail(e, RCA_ERR_ARG, "chain plan: step %d direction %d", i, d);
            int from = d == RCA_DIR_OUT ? e->et_src[t] : e->et_dst[t], to = d == RCA_DIR_OUT ? e->et_dst[t] : e->et_src[t];
            if (i > 0 && from != cur) return fail(e, RCA_ERR_ARG, "chain plan: step %d starts at kind %d but step %d ended at kind %d", i, from, i - 1, cur);
            cur = to;
        }
        return RCA_OK;
    }
    if (pl->kind != RCA_PLAN_VARLEN) return fail(e, RCA_ERR_ARG, "plan kind %d unknown", pl->kind);
    if (pl->max_hops < 1 || pl->max_hops > RCA_MAX_HOPS || pl->min_hops < 1 || pl->min_hops > pl->max_hops)
        return fail(e, RCA_ERR_ARG, "varlen plan: hops %d..%d not within 1..%d", pl->min_hops, pl->max_hops, RCA_MAX_HOPS);
    if (pl->direction < RCA_DIR_OUT || pl->direction > RCA_DIR_UND) return fail(e, RCA_ERR_ARG, "varlen plan: direction %d", pl->direction);
    if (pl->dest_kind >= e->n_kinds) return fail(e, RCA_ERR_ARG, "varlen plan: dest_kind %d unknown", pl->dest_kind);
    for (int i = 0; i <= pl->max_hops; i++)
        if (pl->pos_kind[i] >= e->n_kinds) return fail(e, RCA_ERR_ARG, "varlen plan: pos_kind[%d] = %d unknown", i, pl->pos_kind[i]);
    return RCA_OK;
}

static int row_len_of(const rca_plan_t *pl, int s) {
    return pl->with_states == RCA_STATES_FULL ? 3 * s + 3 : pl->with_states == RCA_STATES_MASK ? 2 * s + 2 : 2 * s + 1;
}

// ---- kernel dispatch: one instantiation per (adjacency format, predicate mode, known depth)
typedef void (*level_kernel_t)(const LevelParams);
struct KernelSet { level_kernel_t count, fill, fused; };
template <bool ADJ2, int MODE, int QT>
static KernelSet kernel_set() { return KernelSet{count_level<ADJ2, MODE, QT>, fill_level<ADJ2, MODE, QT>, fused_level<ADJ2, MODE, QT>}; }
template <bool ADJ2>
static KernelSet pick_kernels(int mode, int q) {
    if (mode == kModeChain) return kernel_set<ADJ2, kModeChain, -1>();
    if (mode == kModeRelUnique) return kernel_set<ADJ2, kModeRelUnique, -1>();
    switch (q) {
        case 0: return kernel_set<ADJ2, kModeSimple, 0>();
        case 1: return kernel_set<ADJ2, kModeSimple, 1>();
        case 2: return kernel_set<ADJ2, kModeSimple, 2>();
        case 3: return kernel_set<ADJ2, kModeSimple, 3>();
        default: return kernel_set<ADJ2, kModeSimple, -1>();
    }
}

// unit steps (rca_kernels.cuh): typed step, at most one candidate per path, path so far of 2..6 nodes
static KernelSet pick_unit_kernels(int q, bool wide = false) {
    if (wide) return q == 1 ? KernelSet{count_unit<1, true>, fill_unit<1, true>, nullptr} : KernelSet{count_unit<2, true>, fill_unit<2, true>, nullptr};
    switch (q) {
        case 1: return KernelSet{count_unit<1>, fill_unit<1>, nullptr};
        case 2: return KernelSet{count_unit<2>, fill_unit<2>, nullptr};
        case 3: return KernelSet{count_unit<3>, fill_unit<3>, nullptr};
        case 4: return KernelSet{count_unit<4>, fill_unit<4>, nullptr};
        default: return KernelSet{count_unit<5>, fill_unit<5>, nullptr};
    }
}

// the fill kernel of a level that writes wide entries (variable-length plans, node-simple, levels 1 and 2)
template <bool ADJ2>
static level_kernel_t pick_wide_fill(int q) {
    return q == 0 ? (level_kernel_t)fill_level<ADJ2, kModeSimple, 0, true> : (level_kernel_t)fill_level<ADJ2, kModeSimple, 1, true>;
}

static int max_grid(rca_engine *e, level_kernel_t kernel, size_t dyn_smem) {
    for (const auto &c : e->occ_cache) if (c.kernel == (const void *)kernel && c.dyn == dyn_smem) return c.grid;
    // levels 6 to 8 need more than the 48 KB of shared memory (static tables included) a kernel gets without asking.
    // The opt-in belongs to the kernel, not to this engine or this (kernel, size) pair: one kernel serves several levels
    // with different sizes, and a later, SMALLER request must not lower what an earlier level was granted (it did: a plan
    // that used the kernel at level 6, then one that first used it at level 5, then level 6 again - found by the
    // randomised parity test).  The largest size ever asked for is kept per device and kernel.
    if (dyn_smem + 8192u > 48u * 1024u) {
        static std::mutex mu;
        static std::vector<std::pair<std::pair<int, const void *>, size_t>> granted;
        std::lock_guard<std::mutex> lock(mu);
        size_t *have = nullptr;
        for (auto &g : granted) if (g.first.first == e->device && g.first.second == (const void *)kernel) have = &g.second;
        if (!have) { granted.push_back({{e->device, (const void *)kernel}, 0}); have = &granted.back().second; }
        if (dyn_smem > *have) {
            const cudaError_t rc = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_smem);
            if (rc != cudaSuccess) std::fprintf(stderr, "rca_b200: cudaFuncSetAttribute(%zu) failed: %s\n", dyn_smem, cudaGetErrorString(rc));
            else *have = dyn_smem;
        }
    }
    int per_sm = 0;
    const cudaError_t oc = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kThreads, dyn_smem);
    if (oc != cudaSuccess || per_sm < 1) {
        if (std::getenv("RCA_B200_DEBUG")) std::fprintf(stderr, "rca_b200: occupancy query (%zu B) -> %s, %d per SM\n", dyn_smem, cudaGetErrorString(oc), per_sm);
        cudaGetLastError();
        per_sm = 1;
    }
    e->occ_cache.push_back({(const void *)kernel, dyn_smem, e->sm_count * per_sm});    // asked once per (kernel, shared memory)
    return e->sm_count * per_sm;
}

extern "C" int rca_set_fused_threshold(rca_engine_t *e, int64_t max_seeds) {
    if (!e || max_seeds < 0) return RCA_ERR_ARG;
    e->fused_max_seeds = e->chain_fused_max_seeds = max_seeds;
    e->plan_cache.clear();
    drop_graphs(e);
    return RCA_OK;
}

extern "C" int rca_set_chain_walk_fanout(rca_engine_t *e, int64_t max_rows_per_seed) {
    if (!e || max_rows_per_seed < 0) return RCA_ERR_ARG;
    e->chain_walk_fanout = max_rows_per_seed;
    e->plan_cache.clear();
    drop_graphs(e);
    return RCA_OK;
}

static void prepare_plan_fresh(rca_engine *e, const rca_plan_t *pl, int64_t n_seeds, unsigned char *t, PreparedPlan &pp) {
    pp.pl = pl;
    pp.varlen = pl->kind == RCA_PLAN_VARLEN;
    pp.L = pp.varlen ? pl->max_hops : pl->n_steps;
    pp.mode = !pp.varlen ? kModeChain : (pl->node_simple ? kModeSimple : kModeRelUnique);
    pp.fused = n_seeds <= (pp.varlen ? e->fused_max_seeds : e->chain_fused_max_seeds);
    if (!pp.varlen && pp.fused && e->chain_walk_fanout > 0) {
        // product of the steps' largest row lengths bounds the rows (and the work) of one seed
        long long fan = 1;
        for (int i = 0; i < pl->n_steps && fan <= e->chain_walk_fanout; i++)
            fan *= std::max(1, pl->step_dir[i] == RCA_DIR_OUT ? e->et_max_out[pl->step_etype[i]] : e->et_max_in[pl->step_etype[i]]);
        pp.walk = fan <= e->chain_walk_fanout;
        pp.fan = fan;
    }
    const int L = pp.L;
    std::memset(t, 0, kPtBytes);
    int *ks = reinterpret_cast<int *>(t + kPtKs);
    for (int k = 0; k < kKindPad; k++) ks[k] = k <= e->n_kinds ? e->kind_start[k] : 0x7fffffff;
    if (pp.varlen) {
        unsigned reach[RCA_MAX_KINDS];
        reach_bits(e, pl, reach);
        for (int s = 0; s <= L; s++) {
            unsigned char *row = t + kPtAct + s * RCA_MAX_KINDS;
            for (int k = 0; k < e->n_kinds; k++) {
                const bool dead = pl->kind_excl[k] || (pl->pos_kind[s] >= 0 && pl->pos_kind[s] != k);
                unsigned a = dead ? kActDead : 0u;
                if (!dead && s < L && ((reach[k] >> (L - s - 1)) & 1u)) a |= kActCont;
                if (!dead && s >= pl->min_hops && (pl->dest_kind < 0 || pl->dest_kind == k)) a |= kActEmit;
                if (pl->kind_inter[k]) a |= kActInter;
                row[k] = (unsigned char)a;
            }
        }
    }
    std::memcpy(t + kPtLut, e->kind_lut.data(), kLutEntries);
    pp.arr_prune = pp.varlen && arrival_reach(e, pl, L, t + kPtArr);

    // Levels that can run as typed sub-launches: when only one or two (edge type, direction) pairs can
    // emit or continue at a level (always the case at the last emitting level of a query with a
    // destination kind), scanning full adjacency rows is wasted work - the typed-range tables lead
    // straight to the useful relationships.  Result-neutral; split path only.
    pp.subs.assign(L + 1, {});
    pp.level_typed.assign(L + 1, 0);
    pp.level_dead.assign(L + 1, 0);
    if (pp.varlen && !pp.fused && pp.arr_prune && pp.mode == kModeSimple) {
        const unsigned char *arr = t + kPtArr, *act = t + kPtAct;
        const ArrivalModel m{e, pl};
        const int A = 2 * e->n_etypes;
        // forward pass: which arrivals can occur at depth s at all.  Depth 1: any allowed hop leaving a
        // kind the seeds may have; depth s+1: any hop that may follow an arrival that continues at depth s.
        std::vector<char> can(A, 0), nxt(A, 0);
        for (int a = 0; a < A; a++) {
            if (!m.allowed(a)) continue;
            can[a] = !(act[0 * RCA_MAX_KINDS + m.tail(a)] & kActDead);     // row 0 holds the seed rules
        }
        pp.cont_kind.assign(L + 1, std::vector<char>(e->n_kinds, 0));
        for (int s = 1; s <= L; s++) {
            std::fill(nxt.begin(), nxt.end(), 0);
            for (int a = 0; a < A; a++) {
                if (!can[a]) continue;
                const unsigned k = act[s * RCA_MAX_KINDS + m.head(a)];
                if (k & kActDead) continue;
                const bool emit = (k & kActEmit) != 0;
                const bool cont = (k & kActCont) && arr[(size_t)s * kArrBytes + a];
                if (cont) pp.cont_kind[s][m.head(a)] = 1;
                if (emit || cont)
                    pp.subs[s].push_back(Sub{a >> 1, a & 1, (cont ? 1 : 0) | (emit ? 2 : 0), m.head(a) | ((k & kActInter) ? 1 << 16 : 0)});
                if (cont)
                    for (int b = 0; b < A; b++) if (!nxt[b] && m.step_ok(a, b)) nxt[b] = 1;
            }
            can = nxt;
            pp.level_dead[s] = pp.subs[s].empty();
            pp.level_typed[s] = !pp.subs[s].empty() && pp.subs[s].size() <= 2;
        }
        // Wide entries (rca_kernels.cuh): level w (1 or 2) stores whole paths when level w+1 consists of unit steps
        // only, nothing continues from it (so nothing ever walks up through level w), and level w itself is ONE
        // launch of the general kernels (its entry count is final when its fill starts).
        if (e->unit_steps && e->wide_entries && !pl->count_only) {
            for (int w = 1; w <= 2 && w < L; w++) {
                const int c = w + 1;
                if (!pp.level_typed[c]) continue;
                bool ok = true;
                for (const Sub &u : pp.subs[c]) ok &= !(u.fl & 1) && (u.d ? e->unit_in_base : e->unit_out_base)[u.t] >= 0;
                bool one_launch = !pp.level_typed[w] || pp.subs[w].size() == 1;
                bool w_unit = false;                  // level w itself a unit level: its kernels do not write wide entries
                if (pp.level_typed[w] && w >= 2)
                    for (const Sub &u : pp.subs[w]) w_unit |= (u.d ? e->unit_in_base : e->unit_out_base)[u.t] >= 0;
                if (ok && one_launch && !w_unit && !pp.level_dead[w]) pp.wide_level = w;
            }
            if (pp.wide_level >= 0 && std::getenv("RCA_B200_DEBUG"))
                std::fprintf(stderr, "rca_b200: plan of %d hops stores level %d as wide entries\n", L, pp.wide_level);
        }
    }
}

// prepare_plan_fresh through a small cache keyed by the plan itself and the one thing the preparation takes
// from the seed count (which family of kernels runs); the knobs that change a preparation clear the cache
static void prepare_plan(rca_engine *e, const rca_plan_t *pl, int64_t n_seeds, unsigned char *t, PreparedPlan &pp) {
    const bool varlen = pl->kind == RCA_PLAN_VARLEN;
    const bool fused = n_seeds <= (varlen ? e->fused_max_seeds : e->chain_fused_max_seeds);
    for (auto &c : e->plan_cache) {
        if (c.fused != fused || std::memcmp(&c.plan, pl, sizeof *pl) != 0) continue;
        pp = c.pp;
        pp.pl = pl;
        std::memcpy(t, c.table.data(), kPtBytes);
        return;
    }
    prepare_plan_fresh(e, pl, n_seeds, t, pp);
    rca_engine::CachedPlan c;
    c.plan = *pl;
    c.fused = fused;
    c.pp = pp;
    c.table.assign(t, t + kPtBytes);
    if (e->plan_cache.size() < 32) e->plan_cache.push_back(std::move(c));
    else e->plan_cache[e->plan_cache_next++ % 32] = std::move(c);
}

// Device buffers one plan runs in.
struct RunBufs {
    const int *seeds; const long long *ts; long long n_seeds;
    Ctl *ctl; const unsigned char *ptab; unsigned epoch; unsigned long long groups;
    unsigned long long tstate_at = 0;            // chain_walk / chain_line: first tile-state word of this plan
    unsigned long long ocap = 0;                 // output words this plan may use (0: the whole arena)
};

// Parameters of a one-launch chain (chain_walk / chain_line / chain_batch).
static void chain_params(rca_engine *e, const PreparedPlan &pp, const RunBufs &b, ChainParams &c, bool &line) {
    const rca_plan_t *pl = pp.pl;
    const int L = pp.L;
    std::memset(&c, 0, sizeof c);
    c.L = L;
    for (int i = 0; i < L; i++) {
        const int t = pl->step_etype[i];
        const bool out = pl->step_dir[i] == RCA_DIR_OUT;
        const int rowk = out ? e->et_src[t] : e->et_dst[t];
        c.rng[i] = out ? e->typed_out + e->typed_out_base[t] : e->typed_in + e->typed_in_base[t];
        const long long ub = e->unit_steps ? (out ? e->unit_out_base : e->unit_in_base)[t] : -1;
        c.unit[i] = ub >= 0 ? e->unit_next + ub : nullptr;
        c.row_lo[i] = e->kind_start[rowk];
        c.row_hi[i] = e->kind_start[rowk + 1];
        c.against[i] = out ? 0 : 1;
    }
    c.adj1 = e->out_dst; c.adj2 = e->in_adj; c.n_nodes = e->n_nodes;
    c.with_states = pl->with_states; c.count_only = pl->count_only; c.row_len = row_len_of(pl, L);
    c.epoch = b.epoch; c.seeds = b.seeds; c.seed_ts = b.ts; c.n_seeds = b.n_seeds;
    c.out = e->out; c.ocap = b.ocap ? b.ocap : e->ocap; c.ctl = b.ctl; c.tstate = e->tstate + b.tstate_at;
    c.st_head = e->st_head; c.st_more = e->st_more; c.n_head = e->n_head; c.st_off = e->st_off; c.st_rec = e->st_rec;
    line = true;                                      // all steps before the last are unit steps: no search needed
    for (int i = 0; i + 1 < L; i++) line &= c.unit[i] != nullptr;
}

// Timing events: inside a stream capture they become event-record nodes that later graph launches re-record
static cudaError_t record_timing(rca_engine *e, cudaEvent_t ev, cudaStream_t st) {
    return e->capturing ? cudaEventRecordWithFlags(ev, st, cudaEventRecordExternal) : cudaEventRecord(ev, st);
}

static void drop_graphs(rca_engine *e) {
    for (auto &g : e->many_graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
    e->many_graphs.clear();
    e->many_seen.clear();
}

// Enqueue every level of a prepared plan on the engine's stream.  ev_lv (nullable): one event per level.
static int launch_levels(rca_engine *e, const PreparedPlan &pp, const RunBufs &b, cudaEvent_t *ev_lv, int *launches) {
    const rca_plan_t *pl = pp.pl;
    const int L = pp.L;
    if (pp.walk) {                                    // the whole chain in one launch
        ChainParams c;
        bool line;
        chain_params(e, pp, b, c, line);
        const long long blocks = std::max<long long>(1, (b.n_seeds + kThreads - 1) / kThreads);
        if (line) chain_line<<<(unsigned)std::min<long long>(blocks, 148LL * 8), kThreads, 0, e->stream>>>(c);
        else chain_walk<<<(unsigned)std::min<long long>(blocks, 148LL * 8), kThreads, 0, e->stream>>>(c);
        ++*launches;
        CK(cudaGetLastError());
        if (ev_lv) CK(record_timing(e, ev_lv[L], e->stream));    // one launch: the earlier levels have no time of their own
        return RCA_OK;
    }
    for (int s = 1; s <= L; s++) {
        LevelParams p;
        std::memset(&p, 0, sizeof p);
        int dir;
        if (pp.varlen) {
            dir = pl->direction;
            p.rng = dir == RCA_DIR_OUT ? e->out_rng : dir == RCA_DIR_IN ? e->in_rng : e->und_rng;
            p.entity_only = pl->entity_only;
        } else {
            const int t = pl->step_etype[s - 1];
            dir = pl->step_dir[s - 1];
            const int rowk = dir == RCA_DIR_OUT ? e->et_src[t] : e->et_dst[t];
            p.typed = 1;
            p.rng = (dir == RCA_DIR_OUT ? e->typed_out + e->typed_out_base[t] : e->typed_in + e->typed_in_base[t]);
            p.row_lo = e->kind_start[rowk];
            p.row_hi = e->kind_start[rowk + 1];
        }
        p.adj1 = e->out_dst;
        p.adj2 = dir == RCA_DIR_IN ? e->in_adj : e->und_adj;
        p.n_nodes = e->n_nodes;
        p.s = s;
        p.chain_flag = s < L ? 1 : 2;
        p.need_inter = pp.varlen && pl->has_inter;
        p.with_states = pl->with_states;
        p.shortest_only = pp.varlen && pl->shortest_only;
        p.count_only = pl->count_only;
        p.row_len = row_len_of(pl, s);
        for (int j = 0; j <= RCA_MAX_HOPS; j++) p.row_len_lv[j] = row_len_of(pl, j);
        p.epoch = b.epoch;
        p.lut_shift = e->lut_shift;
        p.seeds = b.seeds; p.seed_ts = b.ts;
        p.frontier = e->frontier; p.fcap = e->fcap;
        p.out = e->out; p.ocap = b.ocap ? b.ocap : e->ocap;
        p.ctl = b.ctl; p.tstate = e->tstate;
        p.tcount = e->tcount; p.gsum = e->gsum; p.gscan = e->gscan;
        p.ptab = b.ptab;
        p.rel_et = dir == RCA_DIR_OUT ? e->rel_et : dir == RCA_DIR_IN ? e->in_et : e->und_et;   // per slot of the array walked
        p.arr_prune = pp.arr_prune;
        p.st_head = e->st_head; p.st_more = e->st_more; p.n_head = e->n_head; p.st_off = e->st_off; p.st_rec = e->st_rec;
        const size_t dyn = (size_t)kWarps * warp_smem_ints(s) * sizeof(int);
        const long long first = std::max<long long>(1, (b.n_seeds + kTile * kWarps - 1) / (kTile * kWarps));
        // count-only plan whose last level is one unit step that emits: level L-1 takes that step while it counts
        // (rca_kernels.cuh: LevelParams::la_n) - no partial path of level L-1 is stored, level L is not launched
        // does sub-launch j of level lv run through count_unit / fill_unit ?
        auto takes_unit = [&](int lv, int j) {
            return !pp.fused && pp.level_typed[lv] && e->unit_steps && pp.mode == kModeSimple && lv >= 2 && lv <= 6 &&
                   (pp.subs[lv][j].d ? e->unit_in_base : e->unit_out_base)[pp.subs[lv][j].t] >= 0;
        };
        bool merged_last = !pp.fused && pl->count_only && !pl->shortest_only && e->unit_steps && pp.mode == kModeSimple && L >= 2 &&
                           pp.level_typed[L] && pp.subs[L].size() == 1 && (pp.subs[L][0].fl & 2) &&
                           (pp.subs[L][0].d ? e->unit_in_base : e->unit_out_base)[pp.subs[L][0].t] >= 0;
        if (merged_last && pp.level_typed[L - 1])     // ... unless level L-1 is itself a unit step (no fan-out to save on)
            for (size_t j = 0; j < pp.subs[L - 1].size(); j++) merged_last &= !takes_unit(L - 1, (int)j);
        if (merged_last && s == L) {
            if (ev_lv) CK(record_timing(e, ev_lv[s], e->stream));
            continue;
        }
        if (merged_last && s == L - 1) {
            const Sub &u = pp.subs[L][0];
            const int rowk = u.d ? e->et_dst[u.t] : e->et_src[u.t];
            p.la_n = 1;
            p.la_lo = e->kind_start[rowk];
            p.la_hi = e->kind_start[rowk + 1];
            p.la_next = e->unit_next + (u.d ? e->unit_in_base : e->unit_out_base)[u.t];
        }
        if (pp.fused) {
            const KernelSet ks = dir == RCA_DIR_OUT ? pick_kernels<false>(pp.mode, s - 1) : pick_kernels<true>(pp.mode, s - 1);
            int grid = max_grid(e, ks.fused, dyn);
            if (s == 1) grid = (int)std::min<long long>(grid, first);
            ks.fused<<<grid, kThreads, dyn, e->stream>>>(p);
            ++*launches;
        } else if (pp.level_dead[s]) {
            // no (edge type, direction) can emit or continue here: the level and all later ones are empty
        } else {
            const int n_sub = pp.level_typed[s] ? (int)pp.subs[s].size() : 1;
            for (int j = 0; j < n_sub; j++) {
                int sdir = dir;
                if (pp.level_typed[s]) {              // one typed step: ranges of (type, direction) over its near kind
                    const int t = pp.subs[s][j].t;
                    sdir = pp.subs[s][j].d ? RCA_DIR_IN : RCA_DIR_OUT;
                    const int rowk = sdir == RCA_DIR_OUT ? e->et_src[t] : e->et_dst[t];
                    p.typed = 1;
                    p.rng = (sdir == RCA_DIR_OUT ? e->typed_out + e->typed_out_base[t] : e->typed_in + e->typed_in_base[t]);
                    p.row_lo = e->kind_start[rowk];
                    p.row_hi = e->kind_start[rowk + 1];
                    p.adj2 = e->in_adj;
                    p.rel_et = sdir == RCA_DIR_OUT ? e->rel_et : e->in_et;
                }
                p.sub = j;
                KernelSet ks = sdir == RCA_DIR_OUT ? pick_kernels<false>(pp.mode, s - 1) : pick_kernels<true>(pp.mode, s - 1);
                size_t dyn_c = dyn, dyn_f = dyn;
                p.wide_out = pp.wide_level == s;
                p.wide_in = pp.wide_level == s - 1 && s >= 2;
                if (p.wide_out) {
                    ks.fill = sdir == RCA_DIR_OUT ? pick_wide_fill<false>(s - 1) : pick_wide_fill<true>(s - 1);
                    dyn_f = (size_t)kWarps * warp_smem_ints(s, 1) * sizeof(int);
                }
                if (takes_unit(s, j)) {
                    ks = pick_unit_kernels(s - 1, p.wide_in != 0);
                    p.unit_next = e->unit_next + (sdir == RCA_DIR_OUT ? e->unit_out_base : e->unit_in_base)[pp.subs[s][j].t];
                    p.unit_fl = pp.subs[s][j].fl;
                    p.unit_aux = pp.subs[s][j].aux;
                    {   // can a node at positions 0 .. s-2 be of the step's far kind ?  Position 0 is the seed (its kind must
                        // be declared), positions 1 .. s-2 are nodes paths continued through at those levels.
                        const int far = pp.subs[s][j].aux & 0xffff;
                        bool clear = pl->pos_kind[0] >= 0 && pl->pos_kind[0] != far && !pp.cont_kind.empty();
                        for (int d = 1; d <= s - 2 && clear; d++) clear = !pp.cont_kind[d][far];
                        if (clear) p.unit_fl |= 4;
                    }
                    dyn_c = 0;
                    dyn_f = (size_t)kWarps * kTile * p.row_len * sizeof(int);
                }
                if (j) CK(cudaMemsetAsync(e->gsum, 0, b.groups * sizeof(ulonglong2), e->stream));
                int grid = max_grid(e, ks.count, dyn_c);
                if (s == 1) grid = (int)std::min<long long>(grid, first);
                ks.count<<<grid, kThreads, dyn_c, e->stream>>>(p);
                scan_groups<<<1, 1024, 0, e->stream>>>(p);
                if (pl->count_only && (s == L || p.la_n)) {   // nothing is stored: no row, and no path a later level would read
                    *launches += 2;
                    continue;
                }
                grid = max_grid(e, ks.fill, dyn_f);
                if (s == 1) grid = (int)std::min<long long>(grid, first);
                ks.fill<<<grid, kThreads, dyn_f, e->stream>>>(p);
                *launches += 3;
            }
        }
        {
            const cudaError_t le = cudaGetLastError();
            if (le != cudaSuccess)
                return fail(e, RCA_ERR_CUDA, "launch of level %d of %d failed: %s (%s, %d dynamic shared bytes per CTA)", s, L,
                            cudaGetErrorString(le), pp.fused ? "fused" : "count/scan/fill", (int)dyn);
        }
        if (ev_lv) CK(record_timing(e, ev_lv[s], e->stream));
    }
    return RCA_OK;
}

// Tile-slot arrays for a run whose largest seed set is n_seeds (sizes follow the frontier capacity).
static int ensure_tile_arrays(rca_engine *e, long long n_seeds, bool any_fused, bool any_split, unsigned long long *groups_out,
                              unsigned long long extra_slots = 0) {
    const unsigned long long slots = ((unsigned long long)n_seeds + e->fcap) / kTile + (unsigned long long)kGroup * (kLv + 2) + extra_slots;
    const unsigned long long groups = slots / kGroup + kLv + 2;
    int rc;
    if (any_fused && slots > e->tcap) {               // a fresh array starts from zeros and epoch 1
        rc = grow(e, &e->tstate, &e->tcap, slots);
        if (rc) return rc;
        CK(cudaMemsetAsync(e->tstate, 0, e->tcap * sizeof(ulonglong2), e->stream));
        e->epoch = 0;
    }
    if (any_split) {
        rc = grow(e, &e->tcount, &e->ccap, slots);
        if (rc) return rc;
        if (groups > e->gcap) {
            unsigned long long c1 = e->gcap, c2 = e->gcap;
            rc = grow(e, &e->gsum, &c1, groups);
            if (rc) return rc;
            rc = grow(e, &e->gscan, &c2, groups);
            if (rc) return rc;
            e->gcap = groups;
        }
    }
    *groups_out = groups;
    return RCA_OK;
}

// next tile-state tag; when the 30-bit tag wraps, the array is cleared and the count restarts
static int next_epoch(rca_engine *e, unsigned *out) {
    if ((++e->epoch & 0x3fffffffu) == 0) {
        if (e->tstate) CK(cudaMemsetAsync(e->tstate, 0, e->tcap * sizeof(ulonglong2), e->stream));
        e->epoch = 1;
    }
    *out = e->epoch;
    return RCA_OK;
}

static void fill_stats(rca_stats_t *stats, const rca_plan_t *pl, int L, long long n_seeds, const Ctl *c, float ms, int launches,
                       int retries) {
    std::memset(stats, 0, sizeof *stats);
    stats->frontier[0] = n_seeds;
    for (int s = 1; s <= L; s++) {
        stats->frontier[s] = (int64_t)c->F[s];
        stats->scanned[s] = (int64_t)c->E[s];
        stats->emitted[s] = (int64_t)c->N[s];
        stats->total += stats->emitted[s];
    }
    for (int s = 0; s <= L; s++) stats->row_len[s] = row_len_of(pl, s);
    stats->kernel_ms = ms;
    stats->launches = launches;
    stats->retries = retries;
}

static void publish_result(rca_engine *e, PlanResult &r, const rca_plan_t *pl, int L, const Ctl *c) {
    r.levels = L;
    r.valid = !pl->count_only;
    unsigned long long obase = c->out_base;
    for (int s = 0; s <= RCA_MAX_HOPS; s++) { r.rows[s] = 0; r.row_len[s] = row_len_of(pl, s); r.base[s] = 0; }
    for (int s = 1; s <= L; s++) {
        r.rows[s] = (long long)c->N[s];
        r.base[s] = obase;
        obase += c->N[s] * (unsigned long long)row_len_of(pl, s);
    }
}

static int stage_seeds(rca_engine *e, const int32_t *seeds, const int64_t *seed_ts, long long n, bool want_ts, long long at,
                       long long total) {
    unsigned long long cap = (unsigned long long)e->seeds_cap;
    int rc = grow(e, &e->seeds_buf, &cap, (unsigned long long)total);
    e->seeds_cap = (long long)cap;
    if (rc) return rc;
    if (n) CK(cudaMemcpyAsync(e->seeds_buf + at, seeds, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, e->stream));
    if (want_ts) {
        cap = (unsigned long long)e->ts_cap;
        rc = grow(e, &e->ts_buf, &cap, (unsigned long long)total);
        e->ts_cap = (long long)cap;
        if (rc) return rc;
        if (n) CK(cudaMemcpyAsync(e->ts_buf + at, seed_ts, (size_t)n * sizeof(long long), cudaMemcpyHostToDevice, e->stream));
    }
    return RCA_OK;
}

// With double buffering on, a submission writes into the arena the previous one did not use, after the
// asynchronous copies that still read that arena (issued two submissions ago) are through.
static int begin_submission(rca_engine *e) {
    if (!e->double_buffer) return RCA_OK;
    std::swap(e->out, e->out_alt);
    std::swap(e->ocap, e->ocap_alt);
    e->arena ^= 1;
    if (e->copies_pending[e->arena]) {
        CK(cudaStreamWaitEvent(e->stream, e->ev_copied[e->arena], 0));
        e->copies_pending[e->arena] = false;
    }
    return RCA_OK;
}

extern "C" int rca_set_double_buffer(rca_engine_t *e, int on) {
    if (!e) return RCA_ERR_ARG;
    CK(cudaSetDevice(e->device));
    if (on && !e->copy_stream) {
        CK(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
        for (auto &ev : e->ev_copied) CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    }
    if (!on && e->copy_stream) CK(cudaStreamSynchronize(e->copy_stream));
    e->double_buffer = on != 0;
    return RCA_OK;
}

extern "C" int rca_expand(rca_engine_t *e, const rca_plan_t *pl, const int32_t *seeds, int64_t n_seeds,
                          const int64_t *seed_ts, int seeds_on_device, rca_stats_t *stats) {
    if (!e) return RCA_ERR_ARG;
    if (!pl || n_seeds < 0 || (n_seeds > 0 && !seeds)) return fail(e, RCA_ERR_ARG, "rca_expand: null plan or seeds");
    if (n_seeds > 0x7fffffffLL) return fail(e, RCA_ERR_ARG, "rca_expand: more than 2^31-1 seeds in one call");
    if (pl->with_states && n_seeds > 0 && !seed_ts) return fail(e, RCA_ERR_ARG, "rca_expand: with_states needs seed_ts");
    int rc = validate_plan(e, pl);
    if (rc) return rc;
    CK(cudaSetDevice(e->device));
    rc = begin_submission(e);
    if (rc) return rc;
    e->results.assign(1, PlanResult());
    e->selected = 0;
    // few host seeds: everything the kernels read from the host goes up in one copy (layout: table | control block |
    // timestamps | seeds); otherwise table, control block and seed buffers travel separately as before
    constexpr long long kSmallSeeds = 4096;
    constexpr size_t kSmallCtl = ((size_t)kPtBytes + 255) & ~(size_t)255, kSmallTs = kSmallCtl + 512;
    static_assert(sizeof(Ctl) <= 512, "control block does not fit its slot");
    const bool small = !seeds_on_device && n_seeds <= kSmallSeeds;
    if (small && !e->small_dev) {
        CK(cudaMalloc((void **)&e->small_dev, kSmallTs + (size_t)kSmallSeeds * 12));
        CK(cudaMallocHost((void **)&e->small_host, kSmallTs + (size_t)kSmallSeeds * 12));
    }
    PreparedPlan pp;
    prepare_plan(e, pl, n_seeds, small ? e->small_host : e->h_ptab, pp);
    const int L = pp.L;

    const int *d_seeds = seeds;
    const long long *d_ts = (const long long *)seed_ts;
    Ctl *d_ctl = e->ctl;
    const unsigned char *d_ptab = e->ptab;
    size_t small_bytes = 0;
    if (small) {
        const size_t off_seeds = kSmallTs + (size_t)n_seeds * 8;
        if (pl->with_states && n_seeds) std::memcpy(e->small_host + kSmallTs, seed_ts, (size_t)n_seeds * 8);
        if (n_seeds) std::memcpy(e->small_host + off_seeds, seeds, (size_t)n_seeds * 4);
        d_ts = pl->with_states ? (const long long *)(e->small_dev + kSmallTs) : nullptr;
        d_seeds = (const int *)(e->small_dev + off_seeds);
        d_ctl = (Ctl *)(e->small_dev + kSmallCtl);
        d_ptab = e->small_dev;
        small_bytes = off_seeds + (size_t)n_seeds * 4;
    } else {
        if (!seeds_on_device && n_seeds) {
            rc = stage_seeds(e, seeds, seed_ts, n_seeds, pl->with_states != 0, 0, n_seeds);
            if (rc) return rc;
            d_seeds = e->seeds_buf;
            if (pl->with_states) d_ts = e->ts_buf;
        }
        CK(cudaMemcpyAsync(e->ptab, e->h_ptab, kPtBytes, cudaMemcpyHostToDevice, e->stream));
    }

    // scratch arenas: keep what earlier calls grew to, else a guess from the seed count
    if (!e->frontier) { rc = grow(e, &e->frontier, &e->fcap, std::max<unsigned long long>(8ull * n_seeds, 1ull << 20)); if (rc) return rc; }
    if (!e->out) { rc = grow(e, &e->out, &e->ocap, std::max<unsigned long long>(64ull * n_seeds, 1ull << 22)); if (rc) return rc; }

    int retries = 0, launches = 0;
    float ms = 0.f;
    while (true) {
        RunBufs b;
        b.seeds = d_seeds; b.ts = d_ts; b.n_seeds = n_seeds; b.ctl = d_ctl; b.ptab = d_ptab; b.epoch = 0;
        rc = ensure_tile_arrays(e, n_seeds, pp.fused, !pp.fused, &b.groups);
        if (rc) return rc;
        if (pp.fused) { rc = next_epoch(e, &b.epoch); if (rc) return rc; }
        launches = 0;
        std::memset(e->h_ctl, 0, sizeof(Ctl));
        e->h_ctl->F[0] = (unsigned long long)n_seeds;
        CK(cudaEventRecord(e->ev0, e->stream));
        if (small && retries == 0) {                   // table, control block, timestamps and seeds: one copy
            std::memcpy(e->small_host + kSmallCtl, e->h_ctl, sizeof(Ctl));
            CK(cudaMemcpyAsync(e->small_dev, e->small_host, small_bytes, cudaMemcpyHostToDevice, e->stream));
        } else {
            CK(cudaMemcpyAsync(d_ctl, e->h_ctl, sizeof(Ctl), cudaMemcpyHostToDevice, e->stream));
        }
        if (!pp.fused) CK(cudaMemsetAsync(e->gsum, 0, b.groups * sizeof(ulonglong2), e->stream));
        CK(cudaEventRecord(e->ev_lv[0], e->stream));
        rc = launch_levels(e, pp, b, e->ev_lv, &launches);
        if (rc) return rc;
        CK(cudaEventRecord(e->ev1, e->stream));
        CK(cudaMemcpyAsync(e->h_ctl, d_ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
        CK(cudaEventElapsedTime(&ms, e->ev0, e->ev1));
        const unsigned ov = e->h_ctl->overflow;
        if (!ov) break;
        if (++retries > 12) return fail(e, RCA_ERR_NOMEM, "rca_expand: scratch arenas still too small after %d retries", retries);
        // counters are exact up to and including the first level that overflowed: grow to at least that
        unsigned long long fneed = 0, oneed = 0;
        for (int s = 1; s <= L; s++) {
            fneed += s == pp.wide_level ? 2 * (e->h_ctl->F[s] + 32) : e->h_ctl->F[s];      // wide entries: two planes
            oneed += e->h_ctl->N[s] * (unsigned long long)row_len_of(pl, s);
        }
        if (ov & 1u) { rc = grow(e, &e->frontier, &e->fcap, std::max(e->fcap * 2, fneed + fneed / 4)); if (rc) return rc; }
        if (ov & 2u) { rc = grow(e, &e->out, &e->ocap, std::max(e->ocap * 2, oneed + oneed / 4)); if (rc) return rc; }
    }
    publish_result(e, e->results[0], pl, L, e->h_ctl);
    if (stats) {
        fill_stats(stats, pl, L, n_seeds, e->h_ctl, ms, launches, retries);
        for (int s = pp.walk ? L : 1; s <= L; s++) {       // a one-launch chain has one time: the last level's
            float t = 0.f;
            if (cudaEventElapsedTime(&t, e->ev_lv[pp.walk ? 0 : s - 1], e->ev_lv[s]) == cudaSuccess) stats->level_ms[s] = t;
        }
    }
    return RCA_OK;
}

// Several plans in one submission: everything is enqueued back to back on the stream - one table
// upload, one control-block upload, one download, ONE synchronisation for the whole set.  Each
// plan's rows follow the previous plan's in the output arena (the base travels on the device).
extern "C" int rca_expand_many(rca_engine_t *e, const rca_plan_t *plans, int32_t n_plans, const int32_t *const *seeds,
                               const int64_t *n_seeds, const int64_t *const *seed_ts, int seeds_on_device,
                               rca_stats_t *stats) {
    if (!e) return RCA_ERR_ARG;
    if (n_plans < 0 || (n_plans > 0 && (!plans || !seeds || !n_seeds))) return fail(e, RCA_ERR_ARG, "rca_expand_many: null argument");
    e->results.clear();
    e->selected = 0;
    if (n_plans == 0) return RCA_OK;
    long long total_seeds = 0, max_seeds = 0;
    bool any_ts = false;
    for (int k = 0; k < n_plans; k++) {
        if (n_seeds[k] < 0 || n_seeds[k] > 0x7fffffffLL || (n_seeds[k] > 0 && !seeds[k]))
            return fail(e, RCA_ERR_ARG, "rca_expand_many: plan %d: bad seeds", k);
        if (plans[k].with_states && n_seeds[k] > 0 && (!seed_ts || !seed_ts[k]))
            return fail(e, RCA_ERR_ARG, "rca_expand_many: plan %d: with_states needs seed_ts", k);
        int rc = validate_plan(e, &plans[k]);
        if (rc) return rc;
        total_seeds += n_seeds[k];
        max_seeds = std::max<long long>(max_seeds, n_seeds[k]);
        any_ts |= plans[k].with_states != 0;
    }
    CK(cudaSetDevice(e->device));
    int rc = begin_submission(e);
    if (rc) return rc;
    // per-plan tables and control blocks, host (pinned) and device
    const size_t pt_stride = ((size_t)kPtBytes + 255) & ~(size_t)255;
    if ((size_t)n_plans > e->many_cap) {
        CK(cudaStreamSynchronize(e->stream));
        if (e->m_ptab) CK(cudaFree(e->m_ptab));
        if (e->m_ctl) CK(cudaFree(e->m_ctl));
        if (e->mh_ptab) CK(cudaFreeHost(e->mh_ptab));
        if (e->mh_ctl) CK(cudaFreeHost(e->mh_ctl));
        e->m_ptab = nullptr; e->m_ctl = nullptr; e->mh_ptab = nullptr; e->mh_ctl = nullptr; e->many_cap = 0;
        const size_t cap = (size_t)n_plans + 8;
        CK(cudaMalloc((void **)&e->m_ptab, cap * pt_stride));
        CK(cudaMalloc((void **)&e->m_ctl, cap * sizeof(Ctl)));
        CK(cudaMallocHost((void **)&e->mh_ptab, cap * pt_stride));
        CK(cudaMallocHost((void **)&e->mh_ctl, cap * sizeof(Ctl)));
        e->many_cap = cap;
    }
    const size_t ev_stride = RCA_MAX_HOPS + 2;        // per plan: start, then one event per level
    while (e->ev_many.size() < (size_t)n_plans * ev_stride) {
        cudaEvent_t ev;
        CK(cudaEventCreate(&ev));
        e->ev_many.push_back(ev);
    }
    std::vector<PreparedPlan> pps((size_t)n_plans);
    bool any_fused = false, any_split = false;
    for (int k = 0; k < n_plans; k++) {
        prepare_plan(e, &plans[k], n_seeds[k], e->mh_ptab + (size_t)k * pt_stride, pps[k]);
        any_fused |= pps[k].fused;
        any_split |= !pps[k].fused;
    }
    // seeds: one staging buffer holds every plan's seeds back to back (the copies themselves are part of the run)
    std::vector<long long> at((size_t)n_plans, 0);
    std::vector<int> staged_by((size_t)n_plans, -1);
    if (!seeds_on_device) {
        long long off = 0;
        for (int k = 0; k < n_plans; k++) {
            int same = -1;                            // plans of a submission often share one seed list: it is copied once
            for (int j = 0; j < k && same < 0; j++) {
                if (seeds[j] != seeds[k] || n_seeds[j] != n_seeds[k]) continue;
                if (plans[k].with_states && !(plans[j].with_states && seed_ts[j] == seed_ts[k])) continue;
                same = j;
            }
            if (same >= 0) { at[k] = at[same]; staged_by[k] = staged_by[same]; continue; }
            at[k] = off;
            staged_by[k] = k;
            off += n_seeds[k];
        }
        unsigned long long cap = (unsigned long long)e->seeds_cap;
        rc = grow(e, &e->seeds_buf, &cap, (unsigned long long)total_seeds);
        e->seeds_cap = (long long)cap;
        if (rc) return rc;
        if (any_ts) {
            cap = (unsigned long long)e->ts_cap;
            rc = grow(e, &e->ts_buf, &cap, (unsigned long long)total_seeds);
            e->ts_cap = (long long)cap;
            if (rc) return rc;
        }
    }
    if (!e->frontier) { rc = grow(e, &e->frontier, &e->fcap, std::max<unsigned long long>(8ull * max_seeds, 1ull << 20)); if (rc) return rc; }
    if (!e->out) { rc = grow(e, &e->out, &e->ocap, std::max<unsigned long long>(64ull * total_seeds, 1ull << 22)); if (rc) return rc; }

    // Plans that run as one chain_walk / chain_line launch have at most (seeds x fan-out) rows: their rows
    // get fixed places at the top of the output arena, nothing of theirs depends on the plans before
    // them, and they run on side streams beside the other plans (a 100 k-seed chain fills a third of the
    // GPU and is latency-bound).  The other plans follow one another from the bottom of the arena as
    // before, the base travelling on the device.
    std::vector<char> aside((size_t)n_plans, 0);
    std::vector<unsigned long long> top_base((size_t)n_plans, 0), tile_at((size_t)n_plans, 0);
    unsigned long long top_words = 0, side_tiles = 0;
    int n_aside = 0;
    for (int k = 0; k < n_plans; k++) {
        if (!pps[k].walk || n_plans < 2) continue;
        aside[k] = 1;
        n_aside++;
        tile_at[k] = side_tiles;
        side_tiles += (unsigned long long)n_seeds[k] / kThreads + 2;
        if (!plans[k].count_only) top_words += (unsigned long long)n_seeds[k] * (unsigned long long)pps[k].fan * row_len_of(&plans[k], pps[k].L);
    }
    if (n_aside) {
        if (e->ocap < top_words + (1ull << 22)) { rc = grow(e, &e->out, &e->ocap, top_words + std::max<unsigned long long>(64ull * total_seeds, 1ull << 22)); if (rc) return rc; }
        const size_t want = std::min<size_t>((size_t)n_aside, 4);
        while (e->side.size() < want) {
            cudaStream_t st;
            int least = 0, greatest = 0;
            CK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
            CK(cudaStreamCreateWithPriority(&st, cudaStreamNonBlocking, least));     // the main stream's kernels go first
            e->side.push_back(st);
            cudaEvent_t ev;
            CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
            e->ev_side.push_back(ev);
        }
        if (!e->ev_fork) CK(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
    }

    int retries = 0;
    std::vector<int> launches((size_t)n_plans, 0), leader((size_t)n_plans, -1);
    // host seed buffers can go into a graph only if they are pinned (a memcpy node reads them again at every replay)
    bool graphable = e->use_graphs;
    if (graphable && !seeds_on_device) {
        for (int k = 0; k < n_plans && graphable; k++) {
            const void *ptrs[2] = {n_seeds[k] ? (const void *)seeds[k] : nullptr,
                                   plans[k].with_states && n_seeds[k] ? (const void *)seed_ts[k] : nullptr};
            for (const void *q : ptrs) {
                if (!q) continue;
                cudaPointerAttributes at_;
                if (cudaPointerGetAttributes(&at_, q) != cudaSuccess || at_.type != cudaMemoryTypeHost) { graphable = false; cudaGetLastError(); }
            }
        }
    }
    while (true) {
        unsigned long long groups = 0;
        rc = ensure_tile_arrays(e, max_seeds, any_fused, any_split, &groups, side_tiles);
        if (rc) return rc;
        const unsigned long long side_tstate = e->tcap - side_tiles;       // the side plans' tile states sit at the end
        {
            unsigned long long at = e->ocap;                                // ... and their rows at the top, plan by plan
            for (int k = 0; k < n_plans; k++) {
                if (!aside[k] || plans[k].count_only) continue;
                at -= (unsigned long long)n_seeds[k] * (unsigned long long)pps[k].fan * row_len_of(&plans[k], pps[k].L);
                top_base[k] = at;
            }
        }
        // control blocks of this run: counters zero, seeds, row bases of the side plans, tile-state tags (the launch
        // parameters carry none - epoch 0 - so that a replayed graph picks up the tag of ITS run)
        for (int k = 0; k < n_plans; k++) {
            std::memset(&e->mh_ctl[k], 0, sizeof(Ctl));
            e->mh_ctl[k].F[0] = (unsigned long long)n_seeds[k];
            if (aside[k]) e->mh_ctl[k].out_base = top_base[k];
            if (pps[k].fused) { rc = next_epoch(e, &e->mh_ctl[k].epoch); if (rc) return rc; }
        }
        // everything the run enqueues, from the uploads to the download of the control blocks
        auto enqueue = [&]() -> int {
            for (int k = 0; k < n_plans; k++) launches[k] = 0, leader[k] = -1;
            CK(cudaMemcpyAsync(e->m_ptab, e->mh_ptab, (size_t)n_plans * pt_stride, cudaMemcpyHostToDevice, e->stream));
            if (!seeds_on_device) {
                for (int k = 0; k < n_plans; k++) {
                    if (staged_by[k] != k || !n_seeds[k]) continue;       // a seed list several plans share is copied once
                    CK(cudaMemcpyAsync(e->seeds_buf + at[k], seeds[k], (size_t)n_seeds[k] * sizeof(int), cudaMemcpyHostToDevice, e->stream));
                    if (plans[k].with_states)
                        CK(cudaMemcpyAsync(e->ts_buf + at[k], seed_ts[k], (size_t)n_seeds[k] * sizeof(long long), cudaMemcpyHostToDevice, e->stream));
                }
            }
            CK(cudaMemcpyAsync(e->m_ctl, e->mh_ctl, (size_t)n_plans * sizeof(Ctl), cudaMemcpyHostToDevice, e->stream));
            cudaStream_t main_stream = e->stream;
            if (n_aside) CK(cudaEventRecord(e->ev_fork, main_stream));          // uploads are done before any side plan starts
            int prev = -1;                                                      // the last plan that runs in the main stream
            auto bufs_of = [&](int k, RunBufs &b) {
                b.seeds = seeds_on_device ? seeds[k] : e->seeds_buf + at[k];
                b.ts = plans[k].with_states ? (seeds_on_device ? (const long long *)seed_ts[k] : e->ts_buf + at[k]) : nullptr;
                b.n_seeds = n_seeds[k]; b.ctl = e->m_ctl + k; b.ptab = e->m_ptab + (size_t)k * pt_stride; b.epoch = 0; b.groups = groups;
            };
            // The side plans go first, kChainBatch of them per launch (chain_batch), all on one side stream: they
            // start on an idle GPU and run beside the first levels of the main-stream plans.  (Enqueued after those,
            // they waited for SM slots held by the main plans' kernels; one launch per plan cost the host more time
            // than the chains take.)  The plans of a batch share its two timing events.
            if (n_aside) {
                cudaStream_t st = e->side[0];
                CK(cudaStreamWaitEvent(st, e->ev_fork, 0));
                int k = 0;
                while (k < n_plans) {
                    ChainBatch cb;
                    std::memset(&cb, 0, sizeof cb);
                    int members[kChainBatch], nb = 0;
                    long long blocks = 1;
                    for (; k < n_plans && nb < kChainBatch; k++) {
                        if (!aside[k]) continue;
                        RunBufs b;
                        bufs_of(k, b);
                        b.tstate_at = side_tstate + tile_at[k];
                        bool line;
                        chain_params(e, pps[k], b, cb.p[nb], line);
                        cb.p[nb].one_ticket = 1;
                        cb.line[nb] = line ? 1 : 0;
                        blocks = std::max<long long>(blocks, (n_seeds[k] + kThreads - 1) / kThreads);
                        members[nb++] = k;
                    }
                    if (!nb) break;
                    cudaEvent_t *ev0 = e->ev_many.data() + (size_t)members[0] * ev_stride;
                    CK(record_timing(e, ev0[0], st));
                    chain_batch<<<dim3((unsigned)blocks, (unsigned)nb), kThreads, 0, st>>>(cb);     // one CTA per tile
                    CK(cudaGetLastError());
                    CK(record_timing(e, ev0[1], st));
                    for (int j = 0; j < nb; j++) { leader[members[j]] = members[0]; launches[members[j]] = j == 0 ? 1 : 0; }
                }
                CK(cudaEventRecord(e->ev_side[0], st));
            }
            for (int k = 0; k < n_plans; k++) {
                if (aside[k]) continue;
                cudaEvent_t *evk = e->ev_many.data() + (size_t)k * ev_stride;
                RunBufs b;
                bufs_of(k, b);
                b.ocap = e->ocap - top_words;
                if (!pps[k].fused) CK(cudaMemsetAsync(e->gsum, 0, groups * sizeof(ulonglong2), e->stream));
                if (prev >= 0) {                          // this plan's rows start where the previous main-stream plan's end
                    RowLens rl;
                    for (int j = 0; j <= RCA_MAX_HOPS; j++) rl.v[j] = plans[prev].count_only ? 0 : row_len_of(&plans[prev], j);
                    chain_plans<<<1, 32, 0, e->stream>>>(e->m_ctl + prev, e->m_ctl + k, rl);
                    CK(cudaGetLastError());
                }
                CK(record_timing(e, evk[0], e->stream));
                int rc2 = launch_levels(e, pps[k], b, evk, &launches[k]);
                if (rc2) return rc2;
                prev = k;
            }
            if (n_aside) CK(cudaStreamWaitEvent(main_stream, e->ev_side[0], 0));    // join
            CK(cudaMemcpyAsync(e->mh_ctl, e->m_ctl, (size_t)n_plans * sizeof(Ctl), cudaMemcpyDeviceToHost, e->stream));
            return RCA_OK;
        };
        // The identity of this submission: the same plans over the same buffers and arenas enqueue the same operations.
        std::vector<unsigned char> key;
        rca_engine::ManyGraph *hit = nullptr;
        if (graphable) {
            auto put = [&key](const void *q, size_t nbytes) { const unsigned char *b = (const unsigned char *)q; key.insert(key.end(), b, b + nbytes); };
            put(&n_plans, sizeof n_plans); put(plans, sizeof(rca_plan_t) * (size_t)n_plans); put(n_seeds, sizeof(int64_t) * (size_t)n_plans);
            put(seeds, sizeof(void *) * (size_t)n_plans);
            for (int k = 0; k < n_plans; k++) { const void *t = plans[k].with_states && seed_ts ? (const void *)seed_ts[k] : nullptr; put(&t, sizeof t); }
            const void *ptrs[] = {e->stream, e->frontier, e->out, e->tstate, e->tcount, e->gsum, e->gscan, e->m_ptab, e->m_ctl, e->mh_ptab, e->mh_ctl,
                                  e->seeds_buf, e->ts_buf, e->ev_many.data()};
            put(ptrs, sizeof ptrs);
            const unsigned long long caps[] = {e->fcap, e->ocap, e->tcap, e->ccap, e->gcap, (unsigned long long)seeds_on_device, groups};
            put(caps, sizeof caps);
            for (auto &g : e->many_graphs) if (g.key == key) hit = &g;
        }
        if (hit) {
            launches = hit->launches; leader = hit->leader;
            hit->stamp = ++e->many_stamp;
            CK(cudaGraphLaunch(hit->exec, e->stream));
        } else if (graphable && !key.empty() && std::find(e->many_seen.begin(), e->many_seen.end(), key) != e->many_seen.end()) {
            // seen twice: capture what an eager run would enqueue, keep the graph, run it
            cudaGraph_t graph = nullptr;
            CK(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
            e->capturing = true;
            rc = enqueue();
            e->capturing = false;
            cudaError_t ce = cudaStreamEndCapture(e->stream, &graph);
            if (rc || ce != cudaSuccess || !graph) {
                if (graph) cudaGraphDestroy(graph);
                cudaGetLastError();
                e->use_graphs = false;                   // this driver / plan mix cannot be captured: stay on the eager path
                graphable = false;
                if (rc) return rc;
                continue;
            }
            rca_engine::ManyGraph g;
            ce = cudaGraphInstantiate(&g.exec, graph, 0);
            cudaGraphDestroy(graph);
            if (ce != cudaSuccess) { cudaGetLastError(); e->use_graphs = false; graphable = false; continue; }
            g.key = key; g.launches = launches; g.leader = leader; g.stamp = ++e->many_stamp;
            if (e->many_graphs.size() >= 6) {            // keep the most recently used few
                size_t old = 0;
                for (size_t j = 1; j < e->many_graphs.size(); j++) if (e->many_graphs[j].stamp < e->many_graphs[old].stamp) old = j;
                cudaGraphExecDestroy(e->many_graphs[old].exec);
                e->many_graphs.erase(e->many_graphs.begin() + (long)old);
            }
            e->many_graphs.push_back(std::move(g));
            CK(cudaGraphLaunch(e->many_graphs.back().exec, e->stream));
        } else {
            rc = enqueue();
            if (rc) return rc;
            if (graphable && !key.empty()) {
                if (e->many_seen.size() >= 4) e->many_seen.erase(e->many_seen.begin());
                e->many_seen.push_back(key);
            }
        }
        CK(cudaStreamSynchronize(e->stream));
        unsigned ov = 0;
        for (int k = 0; k < n_plans; k++) ov |= e->mh_ctl[k].overflow;
        if (!ov) break;
        if (++retries > 12) return fail(e, RCA_ERR_NOMEM, "rca_expand_many: scratch arenas still too small after %d retries", retries);
        unsigned long long fneed = 0, oneed = top_words;
        for (int k = 0; k < n_plans; k++) {
            unsigned long long f = 0;
            const int L = pps[k].L;
            for (int s = 1; s <= L; s++) {
                f += s == pps[k].wide_level ? 2 * (e->mh_ctl[k].F[s] + 32) : e->mh_ctl[k].F[s];
                if (!aside[k]) oneed += e->mh_ctl[k].N[s] * (unsigned long long)row_len_of(&plans[k], s);
            }
            fneed = std::max(fneed, f);
        }
        if (ov & 1u) { rc = grow(e, &e->frontier, &e->fcap, std::max(e->fcap * 2, fneed + fneed / 4)); if (rc) return rc; }
        if (ov & 2u) { rc = grow(e, &e->out, &e->ocap, std::max(e->ocap * 2, oneed + oneed / 4)); if (rc) return rc; }
    }
    e->results.assign((size_t)n_plans, PlanResult());
    for (int k = 0; k < n_plans; k++) {
        publish_result(e, e->results[k], &plans[k], pps[k].L, &e->mh_ctl[k]);
        if (stats) {
            const bool batched = leader[k] >= 0;       // ran inside a chain_batch launch: the batch's two events
            cudaEvent_t *evk = e->ev_many.data() + (size_t)(batched ? leader[k] : k) * ev_stride;
            float ms = 0.f;
            cudaEventElapsedTime(&ms, evk[0], evk[batched ? 1 : pps[k].L]);
            fill_stats(&stats[k], &plans[k], pps[k].L, n_seeds[k], &e->mh_ctl[k], ms, launches[k], retries);
            if (batched) { stats[k].level_ms[pps[k].L] = ms; continue; }
            for (int s = pps[k].walk ? pps[k].L : 1; s <= pps[k].L; s++) {
                float t = 0.f;
                if (cudaEventElapsedTime(&t, evk[pps[k].walk ? 0 : s - 1], evk[s]) == cudaSuccess) stats[k].level_ms[s] = t;
            }
        }
    }
    return RCA_OK;
}

extern "C" int rca_result_select(rca_engine_t *e, int32_t plan_index) {
    if (!e) return RCA_ERR_ARG;
    if (plan_index < 0 || (size_t)plan_index >= e->results.size()) return fail(e, RCA_ERR_ARG, "rca_result_select: no plan %d in the last submission", plan_index);
    e->selected = plan_index;
    return RCA_OK;
}

static const PlanResult *current_result(const rca_engine *e) {
    if (e->selected < 0 || (size_t)e->selected >= e->results.size() || !e->results[e->selected].valid) return nullptr;
    return &e->results[e->selected];
}

extern "C" int rca_result_level(const rca_engine_t *e, int level, const int32_t **dev_rows, int64_t *n_rows, int32_t *row_len) {
    if (!e || level < 0 || level > RCA_MAX_HOPS) return RCA_ERR_ARG;
    const PlanResult *r = current_result(e);
    if (!r) return fail(const_cast<rca_engine *>(e), RCA_ERR_ARG, "rca_result_level: no result held (call rca_expand first)");
    if (dev_rows) *dev_rows = r->rows[level] ? e->out + r->base[level] : nullptr;
    if (n_rows) *n_rows = r->rows[level];
    if (row_len) *row_len = r->row_len[level];
    return RCA_OK;
}

extern "C" int rca_result_copy(rca_engine_t *e, int level, int32_t *host_rows) {
    if (!e || level < 0 || level > RCA_MAX_HOPS) return RCA_ERR_ARG;
    const PlanResult *r = current_result(e);
    if (!r) return fail(e, RCA_ERR_ARG, "rca_result_copy: no result held (call rca_expand first)");
    const size_t words = (size_t)r->rows[level] * (size_t)r->row_len[level];
    if (!words) return RCA_OK;
    if (!host_rows) return fail(e, RCA_ERR_ARG, "rca_result_copy: null buffer");
    CK(cudaSetDevice(e->device));
    CK(cudaMemcpyAsync(host_rows, e->out + r->base[level], words * sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return RCA_OK;
}

extern "C" int rca_result_copy_async(rca_engine_t *e, int level, int32_t *host_rows) {
    if (!e || level < 0 || level > RCA_MAX_HOPS) return RCA_ERR_ARG;
    if (!e->double_buffer) return fail(e, RCA_ERR_ARG, "rca_result_copy_async: switch double buffering on first (rca_set_double_buffer)");
    const PlanResult *r = current_result(e);
    if (!r) return fail(e, RCA_ERR_ARG, "rca_result_copy_async: no result held (call rca_expand first)");
    const size_t words = (size_t)r->rows[level] * (size_t)r->row_len[level];
    if (!words) return RCA_OK;
    if (!host_rows) return fail(e, RCA_ERR_ARG, "rca_result_copy_async: null buffer");
    CK(cudaSetDevice(e->device));
    // the submission that produced the rows ended with a synchronisation: the copy stream may start at once
    CK(cudaMemcpyAsync(host_rows, e->out + r->base[level], words * sizeof(int), cudaMemcpyDeviceToHost, e->copy_stream));
    CK(cudaEventRecord(e->ev_copied[e->arena], e->copy_stream));
    e->copies_pending[e->arena] = true;
    return RCA_OK;
}

extern "C" int rca_result_wait(rca_engine_t *e) {
    if (!e) return RCA_ERR_ARG;
    if (!e->copy_stream) return RCA_OK;
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->copy_stream));
    e->copies_pending[0] = e->copies_pending[1] = false;
    return RCA_OK;
}

// device staging of at least `need` bytes and a pinned host mirror of the same size
static int io_staging(rca_engine *e, size_t need) {
    if (need > e->io_cap) {
        CK(cudaStreamSynchronize(e->stream));
        if (e->io_buf) CK(cudaFree(e->io_buf));
        e->io_buf = nullptr; e->io_cap = 0;
        CK(cudaMalloc(&e->io_buf, need));
        e->io_cap = need;
    }
    if (need > e->io_host_cap) {
        CK(cudaStreamSynchronize(e->stream));
        if (e->io_host) CK(cudaFreeHost(e->io_host));
        e->io_host = nullptr; e->io_host_cap = 0;
        CK(cudaMallocHost(&e->io_host, need));
        e->io_host_cap = need;
    }
    return RCA_OK;
}

static int state_query(rca_engine_t *e, const char *who, const int32_t *nodes, const int64_t *ta, const int64_t *tb, int64_t n,
                       int on_device, int32_t *out_first, int32_t *out_cnt, int32_t *out_ids, int st_lo = 0,
                       int st_hi = 0x7fffffff) {
    if (!e) return RCA_ERR_ARG;
    if (n < 0 || (n > 0 && (!nodes || !ta || !tb || !out_first || !out_cnt))) return fail(e, RCA_ERR_ARG, "%s: null argument", who);
    if (n == 0) return RCA_OK;
    CK(cudaSetDevice(e->device));
    const int *d_nodes = nodes;
    const long long *d_ta = (const long long *)ta, *d_tb = (const long long *)tb;
    int *d_first = out_first, *d_cnt = out_cnt, *d_ids = out_ids;
    if (!on_device) {
        // staging layout: ta[n] | tb[n] (8 B each) | nodes[n] | first[n] | cnt[n] | ids[n][LIMIT]
        const size_t need = (size_t)n * (16 + 4 + 4 + 4 + 4 * RCA_STATE_LIMIT);
        int rc = io_staging(e, need);
        if (rc) return rc;
        char *b = (char *)e->io_buf, *h = (char *)e->io_host;
        d_ta = (long long *)b;
        d_tb = (long long *)(b + 8 * (size_t)n);
        d_nodes = (int *)(b + 16 * (size_t)n);
        d_first = (int *)(b + 20 * (size_t)n);
        d_cnt = (int *)(b + 24 * (size_t)n);
        d_ids = out_ids ? (int *)(b + 28 * (size_t)n) : nullptr;
        std::memcpy(h, ta, 8 * (size_t)n);                     // the three inputs side by side: one copy up
        std::memcpy(h + 8 * (size_t)n, tb, 8 * (size_t)n);
        std::memcpy(h + 16 * (size_t)n, nodes, 4 * (size_t)n);
        CK(cudaMemcpyAsync(b, h, 20 * (size_t)n, cudaMemcpyHostToDevice, e->stream));
    }
    state_lookup<<<(unsigned)((n + 255) / 256), 256, 0, e->stream>>>(e->st_off, e->st_rec, e->n_nodes, d_nodes, d_ta, d_tb, n,
                                                                    st_lo, st_hi, d_first, d_cnt, d_ids);
    CK(cudaGetLastError());
    if (!on_device) {                                         // the three outputs side by side: one copy down
        char *b = (char *)e->io_buf, *h = (char *)e->io_host;
        const size_t out_bytes = (size_t)n * (out_ids ? 8 + 4 * RCA_STATE_LIMIT : 8);
        CK(cudaMemcpyAsync(h + 20 * (size_t)n, b + 20 * (size_t)n, out_bytes, cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
        std::memcpy(out_first, h + 20 * (size_t)n, 4 * (size_t)n);
        std::memcpy(out_cnt, h + 24 * (size_t)n, 4 * (size_t)n);
        if (out_ids) std::memcpy(out_ids, h + 28 * (size_t)n, 4 * (size_t)n * RCA_STATE_LIMIT);
        return RCA_OK;
    }
    CK(cudaStreamSynchronize(e->stream));
    return RCA_OK;
}

extern "C" int rca_state_lookup(rca_engine_t *e, const int32_t *nodes, const int64_t *ts, int64_t n, int on_device,
                                int32_t *out_first, int32_t *out_cnt, int32_t *out_ids) {
    return state_query(e, "rca_state_lookup", nodes, ts, ts, n, on_device, out_first, out_cnt, out_ids);
}

extern "C" int rca_state_window(rca_engine_t *e, const int32_t *nodes, const int64_t *tmin_le, const int64_t *tmax_gt,
                                int64_t n, int on_device, int32_t *out_first, int32_t *out_cnt, int32_t *out_ids) {
    return state_query(e, "rca_state_window", nodes, tmin_le, tmax_gt, n, on_device, out_first, out_cnt, out_ids);
}

extern "C" int rca_state_window_kind(rca_engine_t *e, const int32_t *nodes, const int64_t *tmin_le, const int64_t *tmax_gt,
                                     int64_t n, int on_device, int32_t state_kind, int32_t *out_first, int32_t *out_cnt,
                                     int32_t *out_ids) {
    if (!e) return RCA_ERR_ARG;
    if (state_kind < -1 || state_kind >= e->n_kinds) return fail(e, RCA_ERR_ARG, "rca_state_window_kind: kind %d unknown", state_kind);
    const int lo = state_kind < 0 ? 0 : e->kind_start[state_kind], hi = state_kind < 0 ? 0x7fffffff : e->kind_start[state_kind + 1];
    return state_query(e, "rca_state_window_kind", nodes, tmin_le, tmax_gt, n, on_device, out_first, out_cnt, out_ids, lo, hi);
}

// ------------------------------------------------------------------------------------------------
// Count exchange over peer memory: see peer_exchange in rca_kernels.cuh.
static_assert(sizeof(cudaIpcMemHandle_t) == RCA_PEER_HANDLE_BYTES, "RCA_PEER_HANDLE_BYTES must be the size of cudaIpcMemHandle_t");

extern "C" int rca_peer_handle(rca_engine_t *e, int32_t world, uint8_t *out_handle) {
    if (!e) return RCA_ERR_ARG;
    if (world < 1 || world > 8 || !out_handle) return fail(e, RCA_ERR_ARG, "rca_peer_handle: world %d not in 1..8, or null handle", world);
    if (e->peer_own) return fail(e, RCA_ERR_ARG, "rca_peer_handle: already called on this engine");
    CK(cudaSetDevice(e->device));
    CK(cudaMalloc((void **)&e->peer_own, kPeerBuffers * (size_t)world * sizeof(PeerSlot)));
    CK(cudaMemset(e->peer_own, 0, kPeerBuffers * (size_t)world * sizeof(PeerSlot)));
    CK(cudaMalloc((void **)&e->peer_out, (size_t)world * sizeof(long long)));
    CK(cudaMalloc((void **)&e->peer_err, sizeof(unsigned)));
    CK(cudaMemset(e->peer_err, 0, sizeof(unsigned)));
    CK(cudaMallocHost((void **)&e->peer_out_h, ((size_t)world + 1) * sizeof(long long)));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, e->peer_own));
    std::memcpy(out_handle, &h, sizeof h);
    e->peer_world = world;
    return RCA_OK;
}

extern "C" int rca_peer_connect(rca_engine_t *e, int32_t rank, const uint8_t *handles) {
    if (!e) return RCA_ERR_ARG;
    if (!e->peer_own) return fail(e, RCA_ERR_ARG, "rca_peer_connect: call rca_peer_handle first");
    if (rank < 0 || rank >= e->peer_world || !handles) return fail(e, RCA_ERR_ARG, "rca_peer_connect: rank %d not in 0..%d, or null handles", rank, e->peer_world - 1);
    if (e->peer_rank >= 0) return fail(e, RCA_ERR_ARG, "rca_peer_connect: already connected");
    CK(cudaSetDevice(e->device));
    for (int r = 0; r < e->peer_world; r++) {
        if (r == rank) { e->peer_slots[r] = e->peer_own; continue; }
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handles + (size_t)r * sizeof h, sizeof h);
        void *ptr = nullptr;
        CK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
        e->peer_slots[r] = (PeerSlot *)ptr;
    }
    e->peer_rank = rank;
    return RCA_OK;
}

extern "C" int rca_peer_set_timeout(rca_engine_t *e, int64_t milliseconds) {
    if (!e || milliseconds < 1) return RCA_ERR_ARG;
    e->peer_timeout_ns = (unsigned long long)milliseconds * 1000000ull;
    return RCA_OK;
}

static int peer_launch(rca_engine *e, unsigned long long seq, int publish, int deferred, long long value) {
    PeerParams p;
    std::memset(&p, 0, sizeof p);
    for (int r = 0; r < e->peer_world; r++) {
        p.slots[r] = e->peer_slots[r] + (seq % kPeerBuffers) * (size_t)e->peer_world;
        p.prev[r] = e->peer_slots[r] + ((seq + kPeerBuffers - 1) % kPeerBuffers) * (size_t)e->peer_world;
    }
    p.rank = e->peer_rank; p.world = e->peer_world; p.seq = seq;
    p.publish = publish; p.deferred = deferred; p.host_value = value;
    p.out = e->peer_out; p.err = e->peer_err; p.timeout_ns = e->peer_timeout_ns;
    peer_exchange<<<1, 32, 0, e->stream>>>(p);
    CK(cudaGetLastError());
    return RCA_OK;
}

extern "C" int rca_peer_allgather(rca_engine_t *e, int64_t value, int64_t *out_values) {
    if (!e) return RCA_ERR_ARG;
    if (e->peer_rank < 0) return fail(e, RCA_ERR_ARG, "rca_peer_allgather: not connected (rca_peer_handle, rca_peer_connect)");
    CK(cudaSetDevice(e->device));
    // with a buffer for the result the exchange is complete when the call returns; without one the value is only
    // published (and the PREVIOUS exchange collected): the ranks stay within one exchange of each other
    int rc = peer_launch(e, ++e->peer_seq, 1, out_values ? 0 : 1, value);
    if (rc) return rc;
    e->peer_collected = out_values != nullptr;
    if (out_values) return rca_peer_wait(e, out_values);
    return RCA_OK;
}

extern "C" int rca_peer_wait(rca_engine_t *e, int64_t *out_values) {
    if (!e) return RCA_ERR_ARG;
    if (e->peer_rank < 0 || !out_values) return fail(e, RCA_ERR_ARG, "rca_peer_wait: not connected, or null buffer");
    CK(cudaSetDevice(e->device));
    if (!e->peer_collected && e->peer_seq) {          // the last exchange was only published: collect it now
        int rc = peer_launch(e, e->peer_seq, 0, 0, 0);
        if (rc) return rc;
        e->peer_collected = true;
    }
    CK(cudaMemcpyAsync(e->peer_out_h, e->peer_out, (size_t)e->peer_world * sizeof(long long), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaMemcpyAsync(e->peer_out_h + e->peer_world, e->peer_err, sizeof(unsigned), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    if (*(const unsigned *)(e->peer_out_h + e->peer_world)) {
        // reported once: the flag is cleared so that later exchanges are judged on their own.  The ranks' exchange
        // counts may now differ (a peer that was late, not dead, is one exchange behind): callers should re-create
        // the engines' peer set-up, or stop.
        CK(cudaMemsetAsync(e->peer_err, 0, sizeof(unsigned), e->stream));
        return fail(e, RCA_ERR_CUDA, "rca_peer_allgather: a peer did not publish its value within %.1f s (rca_peer_set_timeout)",
                    (double)e->peer_timeout_ns * 1e-9);
    }
    for (int r = 0; r < e->peer_world; r++) out_values[r] = e->peer_out_h[r];
    return RCA_OK;
}

// ------------------------------------------------------------------------------------------------
// Row gather over peer memory: see push_rows in rca_kernels.cuh.
extern "C" int rca_gather_window(rca_engine_t *e, int64_t words, uint8_t *out_handle) {
    if (!e) return RCA_ERR_ARG;
    if (words < 1 || !out_handle) return fail(e, RCA_ERR_ARG, "rca_gather_window: words %lld, or null handle", (long long)words);
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->stream));
    if (e->gather_dst && !e->gather_dst_mapped) { e->gather_dst = nullptr; e->gather_dst_rank = -1; }    // it was the old window
    if (e->gather_own) { CK(cudaFree(e->gather_own)); e->gather_own = nullptr; e->gather_own_words = 0; }
    CK(cudaMalloc((void **)&e->gather_own, (size_t)words * sizeof(int)));
    e->gather_own_words = (unsigned long long)words;
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, e->gather_own));
    std::memcpy(out_handle, &h, sizeof h);
    return RCA_OK;
}

extern "C" int rca_gather_connect(rca_engine_t *e, int32_t dst_rank, const uint8_t *handle, int64_t words) {
    if (!e) return RCA_ERR_ARG;
    if (e->peer_rank < 0) return fail(e, RCA_ERR_ARG, "rca_gather_connect: set the count exchange up first (rca_peer_handle, rca_peer_connect)");
    if (dst_rank < 0 || dst_rank >= e->peer_world || words < 1) return fail(e, RCA_ERR_ARG, "rca_gather_connect: rank %d not in 0..%d, or no words", dst_rank, e->peer_world - 1);
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->stream));
    if (e->gather_dst_mapped && e->gather_dst) CK(cudaIpcCloseMemHandle(e->gather_dst));
    e->gather_dst = nullptr; e->gather_dst_mapped = false; e->gather_dst_rank = -1;
    if (dst_rank == e->peer_rank) {
        if (!e->gather_own || e->gather_own_words < (unsigned long long)words)
            return fail(e, RCA_ERR_ARG, "rca_gather_connect: this rank is the destination but holds no window of %lld words (rca_gather_window)", (long long)words);
        e->gather_dst = e->gather_own;
    } else {
        if (!handle) return fail(e, RCA_ERR_ARG, "rca_gather_connect: null handle");
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handle, sizeof h);
        void *ptr = nullptr;
        CK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
        e->gather_dst = (int *)ptr;
        e->gather_dst_mapped = true;
    }
    e->gather_dst_words = (unsigned long long)words;
    e->gather_dst_rank = dst_rank;
    if (!e->ev_g0) { CK(cudaEventCreate(&e->ev_g0)); CK(cudaEventCreate(&e->ev_g1)); }
    return RCA_OK;
}

extern "C" int rca_gather_rows(rca_engine_t *e, int level, int64_t window_offset_words, const int32_t **dev_rows,
                               int64_t *n_rows, int64_t *rank_rows, double *push_ms) {
    if (!e || level < 0 || level > RCA_MAX_HOPS || window_offset_words < 0) return RCA_ERR_ARG;
    if (e->gather_dst_rank < 0) return fail(e, RCA_ERR_ARG, "rca_gather_rows: no destination window connected (rca_gather_connect)");
    const PlanResult *r = current_result(e);
    if (!r) return fail(e, RCA_ERR_ARG, "rca_gather_rows: no result held (call rca_expand first)");
    CK(cudaSetDevice(e->device));
    const int world = e->peer_world, rank = e->peer_rank, dst = e->gather_dst_rank;
    // 1. every rank learns every rank's row count (and that all ranks mean rows of the same width)
    int64_t counts[8];
    const long long rl = r->row_len[level];
    int rc = rca_peer_allgather(e, (int64_t)r->rows[level] | ((int64_t)rl << 48), counts);
    if (rc) return rc;
    unsigned long long total = 0, before = 0;
    for (int k = 0; k < world; k++) {
        if ((counts[k] >> 48) != rl) return fail(e, RCA_ERR_ARG, "rca_gather_rows: rank %d holds rows of %lld words, this rank of %lld", k, (long long)(counts[k] >> 48), rl);
        counts[k] &= (1ll << 48) - 1;
        if (k < rank) before += (unsigned long long)counts[k];
        total += (unsigned long long)counts[k];
        if (rank_rows) rank_rows[k] = counts[k];
    }
    if (n_rows) *n_rows = (int64_t)total;
    if (dev_rows) *dev_rows = nullptr;
    const unsigned long long woff = (unsigned long long)window_offset_words;
    if (woff + total * (unsigned long long)rl > e->gather_dst_words)       // every rank sees the same totals: all of them refuse
        return fail(e, RCA_ERR_NOMEM, "rca_gather_rows: %llu rows of %lld words at word %llu do not fit the destination window of %llu words",
                    total, rl, woff, e->gather_dst_words);
    // 2. push this rank's rows to their place in the destination window (rank order = seed order for contiguous shards)
    const unsigned long long words = (unsigned long long)r->rows[level] * (unsigned long long)rl;
    CK(cudaEventRecord(e->ev_g0, e->stream));
    if (words) {
        PushParams pp;
        pp.src = e->out + r->base[level];
        pp.dst = e->gather_dst + woff + before * (unsigned long long)rl;
        pp.words = words;
        const unsigned long long want = (words / 4 + 255) / 256 + 1;
        const unsigned grid = (unsigned)std::min<unsigned long long>(want, (unsigned long long)e->sm_count * 8);
        push_rows<<<grid, 256, 0, e->stream>>>(pp);
        CK(cudaGetLastError());
    }
    CK(cudaEventRecord(e->ev_g1, e->stream));
    // 3. second exchange = completion: its release store follows the push on this stream, so when the destination
    //    has seen every rank's sequence number, every rank's rows are in its window
    int64_t done[8];
    rc = rca_peer_allgather(e, (int64_t)words, done);
    if (rc) return rc;
    if (push_ms) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e->ev_g0, e->ev_g1));
        *push_ms = ms;
    }
    if (rank == dst && dev_rows) *dev_rows = total ? e->gather_own + woff : nullptr;
    return RCA_OK;
}

// ---- message CONTAINS on the device
extern "C" int rca_text_load(rca_engine_t *e, const uint8_t *bytes, const int64_t *off, const int32_t *node_ids, int64_t n_msgs) {
    if (!e) return RCA_ERR_ARG;
    if (n_msgs < 0 || (n_msgs > 0 && (!bytes || !off || !node_ids))) return fail(e, RCA_ERR_ARG, "rca_text_load: null argument");
    if (n_msgs > 0x7ffffff0LL) return fail(e, RCA_ERR_ARG, "rca_text_load: too many messages");
    CK(cudaSetDevice(e->device));
    CK(cudaStreamSynchronize(e->stream));
    if (e->msg_bytes) { CK(cudaFree(e->msg_bytes)); e->msg_bytes = nullptr; }
    if (e->msg_off) { CK(cudaFree(e->msg_off)); e->msg_off = nullptr; }
    e->n_msgs = 0;
    const long long total = n_msgs ? off[n_msgs] : 0;
    CK(upload(&e->msg_bytes, bytes, (size_t)total));
    CK(upload(&e->msg_off, off, (size_t)n_msgs + 1));
    e->msg_node.assign(node_ids, node_ids + n_msgs);
    e->n_msgs = n_msgs;
    return RCA_OK;
}

extern "C" int rca_text_contains(rca_engine_t *e, const uint8_t *pat_bytes, const int64_t *pat_off, int64_t n_pat,
                                 int32_t *out_first_node, int32_t *out_pairs, int64_t pairs_cap, int64_t *n_pairs) {
    if (!e) return RCA_ERR_ARG;
    if (n_pat < 0 || (n_pat > 0 && (!pat_bytes || !pat_off || !out_first_node)) || pairs_cap < 0 || (pairs_cap > 0 && !out_pairs))
        return fail(e, RCA_ERR_ARG, "rca_text_contains: null argument");
    if (n_pairs) *n_pairs = 0;
    if (n_pat == 0) return RCA_OK;
    CK(cudaSetDevice(e->device));
    // distinct lengths, their top powers, and the hash table of the patterns (host, cold path)
    std::vector<int> lens;
    std::vector<unsigned long long> hashes((size_t)n_pat);
    long long pairs_found = 0;
    std::vector<int> first((size_t)n_pat, 0x7fffffff);
    std::vector<int> host_pairs;
    for (int64_t i = 0; i < n_pat; i++) {
        const long long len = pat_off[i + 1] - pat_off[i];
        if (len < 0 || len > 0x7fffffff) return fail(e, RCA_ERR_ARG, "rca_text_contains: bad pattern offsets");
        unsigned long long h = 0;
        for (long long k = 0; k < len; k++) h = h * kRkBase + pat_bytes[pat_off[i] + k];
        hashes[i] = h;
        if (len > 0) lens.push_back((int)len);
    }
    std::sort(lens.begin(), lens.end());
    lens.erase(std::unique(lens.begin(), lens.end()), lens.end());
    std::vector<unsigned long long> pow_top(lens.size());
    for (size_t d = 0; d < lens.size(); d++) {
        unsigned long long v = 1;
        for (int k = 1; k < lens[d]; k++) v *= kRkBase;
        pow_top[d] = v;
    }
    unsigned tsize = 64;
    while (tsize < 4 * (unsigned long long)n_pat) tsize <<= 1;
    std::vector<unsigned long long> tab_hash(tsize, 0);
    std::vector<int> tab_pat(tsize, -1);
    for (int64_t i = 0; i < n_pat; i++) {
        const int len = (int)(pat_off[i + 1] - pat_off[i]);
        if (len == 0) continue;                               // the empty string is contained in every message
        unsigned long long x = hashes[i] ^ ((unsigned long long)len * 0x9e3779b97f4a7c15ull);
        x ^= x >> 29; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 32;
        unsigned sl = (unsigned)x & (tsize - 1);
        while (tab_pat[sl] >= 0) sl = (sl + 1) & (tsize - 1);
        tab_pat[sl] = (int)i;
        tab_hash[sl] = hashes[i];
    }
    // one staging allocation for everything this call needs on the device
    const size_t pbytes = (size_t)pat_off[n_pat];
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    size_t o_pat = 0, o_poff = al(pbytes), o_lens = o_poff + al((n_pat + 1) * 8), o_pow = o_lens + al(lens.size() * 4 + 4),
           o_th = o_pow + al(lens.size() * 8 + 8), o_tp = o_th + al((size_t)tsize * 8), o_first = o_tp + al((size_t)tsize * 4),
           o_np = o_first + al((size_t)n_pat * 4), o_pairs = o_np + 256, need = o_pairs + al((size_t)pairs_cap * 8);
    {
        int rc = io_staging(e, need);
        if (rc) return rc;
    }
    char *b = (char *)e->io_buf, *h = (char *)e->io_host;
    // everything the kernel reads, laid out in the pinned mirror as it lies on the device: ONE copy up
    std::memcpy(h + o_pat, pat_bytes, pbytes);
    std::memcpy(h + o_poff, pat_off, (size_t)(n_pat + 1) * 8);
    if (!lens.empty()) {
        std::memcpy(h + o_lens, lens.data(), lens.size() * 4);
        std::memcpy(h + o_pow, pow_top.data(), lens.size() * 8);
    }
    std::memcpy(h + o_th, tab_hash.data(), (size_t)tsize * 8);
    std::memcpy(h + o_tp, tab_pat.data(), (size_t)tsize * 4);
    std::memcpy(h + o_first, first.data(), (size_t)n_pat * 4);
    std::memset(h + o_np, 0, 8);
    CK(cudaMemcpyAsync(b, h, o_np + 8, cudaMemcpyHostToDevice, e->stream));
    unsigned long long found = 0;
    if (!lens.empty() && e->n_msgs > 0) {
        TextParams tp;
        tp.msg = e->msg_bytes; tp.msg_off = e->msg_off; tp.n_msgs = e->n_msgs;
        tp.pat = (const unsigned char *)(b + o_pat); tp.pat_off = (const long long *)(b + o_poff);
        tp.lens = (const int *)(b + o_lens); tp.pow_top = (const unsigned long long *)(b + o_pow); tp.n_lens = (int)lens.size();
        tp.tab_hash = (const unsigned long long *)(b + o_th); tp.tab_pat = (const int *)(b + o_tp); tp.tab_mask = tsize - 1;
        tp.first = (int *)(b + o_first); tp.pairs = (int2 *)(b + o_pairs); tp.cap = (unsigned long long)pairs_cap;
        tp.n_pairs = (unsigned long long *)(b + o_np);
        const long long threads = e->n_msgs * (long long)lens.size();
        contains_scan<<<(unsigned)((threads + 255) / 256), 256, 0, e->stream>>>(tp);
        CK(cudaGetLastError());
        // the answers - first match per pattern, number of pairs - and the first few pairs come down in ONE copy; only a
        // longer pair list needs a second one
        const unsigned long long eager = std::min<unsigned long long>((unsigned long long)pairs_cap, 64);
        CK(cudaMemcpyAsync(h + o_first, b + o_first, (o_pairs - o_first) + eager * 8, cudaMemcpyDeviceToHost, e->stream));
        CK(cudaStreamSynchronize(e->stream));
        std::memcpy(first.data(), h + o_first, (size_t)n_pat * 4);
        std::memcpy(&found, h + o_np, 8);
        const unsigned long long got = std::min<unsigned long long>(found, (unsigned long long)pairs_cap);
        if (got) {
            if (got > eager) {
                CK(cudaMemcpyAsync(h + o_pairs, b + o_pairs, got * 8, cudaMemcpyDeviceToHost, e->stream));
                CK(cudaStreamSynchronize(e->stream));
            }
            std::memcpy(out_pairs, h + o_pairs, got * 8);
            for (unsigned long long k = 0; k < got; k++) out_pairs[2 * k + 1] = e->msg_node[out_pairs[2 * k + 1]];   // message index -> node id
        }
    } else {
        CK(cudaStreamSynchronize(e->stream));
    }
    pairs_found = (long long)found;
    for (int64_t i = 0; i < n_pat; i++) {
        const bool empty = pat_off[i + 1] == pat_off[i];
        if (empty) out_first_node[i] = e->n_msgs ? e->msg_node[0] : -1;      // callers expand 'every message' themselves
        else out_first_node[i] = first[i] == 0x7fffffff ? -1 : e->msg_node[first[i]];
    }
    if (n_pairs) *n_pairs = pairs_found;
    return RCA_OK;
}

extern "C" int rca_text_match_pairs(rca_engine_t *e, const uint8_t *needle_bytes, const int64_t *needle_off, int64_t n_needles,
                                    const int32_t *msg_index, const int32_t *needle_a, const int32_t *needle_b, int64_t n_pairs,
                                    uint8_t *out_flags) {
    if (!e) return RCA_ERR_ARG;
    if (n_pairs < 0 || n_needles < 0 || (n_pairs > 0 && (!msg_index || !needle_a || !needle_b || !out_flags)) ||
        (n_needles > 0 && (!needle_bytes || !needle_off)))
        return fail(e, RCA_ERR_ARG, "rca_text_match_pairs: null argument");
    if (n_pairs == 0) return RCA_OK;
    if (n_pairs > 0x3fffffffLL) return fail(e, RCA_ERR_ARG, "rca_text_match_pairs: more than 2^30 pairs in one call");
    for (int64_t k = 0; k < n_pairs; k++) {
        if (msg_index[k] < 0 || msg_index[k] >= e->n_msgs)
            return fail(e, RCA_ERR_ARG, "rca_text_match_pairs: pair %lld: message %d is not in the loaded arena (rca_text_load)", (long long)k, msg_index[k]);
        if (needle_a[k] < -1 || needle_a[k] >= n_needles || needle_b[k] < -1 || needle_b[k] >= n_needles)
            return fail(e, RCA_ERR_ARG, "rca_text_match_pairs: pair %lld: needle index out of range", (long long)k);
    }
    CK(cudaSetDevice(e->device));
    const size_t nbytes = n_needles ? (size_t)needle_off[n_needles] : 0;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t o_nb = 0, o_no = al(nbytes + 1), o_hay = o_no + al(((size_t)n_needles + 1) * 8), o_a = o_hay + al((size_t)n_pairs * 4),
                 o_b = o_a + al((size_t)n_pairs * 4), o_out = o_b + al((size_t)n_pairs * 4), need = o_out + al((size_t)n_pairs);
    if (need > e->io_cap) {
        CK(cudaStreamSynchronize(e->stream));
        if (e->io_buf) CK(cudaFree(e->io_buf));
        e->io_buf = nullptr; e->io_cap = 0;
        CK(cudaMalloc(&e->io_buf, need));
        e->io_cap = need;
    }
    char *b = (char *)e->io_buf;
    const long long zero = 0;
    if (nbytes) CK(cudaMemcpyAsync(b + o_nb, needle_bytes, nbytes, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(b + o_no, n_needles ? (const void *)needle_off : (const void *)&zero, ((size_t)n_needles + 1) * 8, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(b + o_hay, msg_index, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(b + o_a, needle_a, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, e->stream));
    CK(cudaMemcpyAsync(b + o_b, needle_b, (size_t)n_pairs * 4, cudaMemcpyHostToDevice, e->stream));
    PairParams pp;
    pp.msg = e->msg_bytes; pp.msg_off = e->msg_off;
    pp.ndl = (const unsigned char *)(b + o_nb); pp.ndl_off = (const long long *)(b + o_no);
    pp.hay = (const int *)(b + o_hay); pp.na = (const int *)(b + o_a); pp.nb = (const int *)(b + o_b);
    pp.n = n_pairs; pp.out = (unsigned char *)(b + o_out);
    match_pairs<<<(unsigned)((n_pairs * 32 + 255) / 256), 256, 0, e->stream>>>(pp);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_flags, b + o_out, (size_t)n_pairs, cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    return RCA_OK;
}

// ================================================================================================
// rca_expand_static: the same expansion for callers that must not allocate.
//
// Who needs this: code that runs where device memory may not be requested - a service that gave
// its whole device-memory budget to a pool at start-up, a handler on a latency path where
// cudaMalloc (which may synchronise the device) is forbidden, a process under a hard memory cap.
// rca_expand grows its scratch arenas with cudaMalloc when a query needs more room; this entry
// point never does.  Its scratch - control block, plan table, tile states and the frontier of
// partial paths - is the fixed array below, which lives in the library image: every GPU that
// loads the library holds RCA_STATIC_SCRATCH_BYTES of it until the process exits.
//
// The array's name is only a host-side handle for the symbol.  Kernels need its address in device
// memory, which cudaGetSymbolAddress returns once per device (with that device current); that
// happens on the first call for a device, see static_scratch_for().
static __device__ __align__(256) unsigned char g_static_scratch[RCA_STATIC_SCRATCH_BYTES];

namespace {
constexpr size_t kSsCtl = 0;                                                    // Ctl
constexpr size_t kSsPtab = 512;                                                 // plan table
constexpr size_t kSsTstate = kSsPtab + ((kPtBytes + 255) / 256) * 256;          // tile-state words
constexpr size_t kSsSlots = (RCA_STATIC_MAX_SEEDS + RCA_STATIC_MAX_FRONTIER) / kTile + (size_t)kGroup * (kLv + 2);
constexpr size_t kSsFrontier = kSsTstate + ((kSsSlots * sizeof(ulonglong2) + 255) / 256) * 256;
constexpr size_t kSsEnd = kSsFrontier + (size_t)RCA_STATIC_MAX_FRONTIER * sizeof(int4);
static_assert(sizeof(Ctl) <= kSsPtab, "control block does not fit its slot");
static_assert(kSsEnd <= RCA_STATIC_SCRATCH_BYTES, "RCA_STATIC_SCRATCH_BYTES is too small for the stated maxima");

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
unsigned char *g_static_ptr[kMaxDevices] = {nullptr};
unsigned g_static_epoch[kMaxDevices] = {0};

// Device address of the scratch array on `device` (which must be the current device); resolved once.
cudaError_t static_scratch_for(int device, unsigned char **out) {
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[device]) {
        void *p = nullptr;
        cudaError_t rc = cudaGetSymbolAddress(&p, g_static_scratch);
        if (rc != cudaSuccess) return rc;
        g_static_ptr[device] = static_cast<unsigned char *>(p);
    }
    *out = g_static_ptr[device];
    return cudaSuccess;
}
}  // namespace

extern "C" int rca_expand_static(rca_engine_t *e, const rca_plan_t *pl, const int32_t *dev_seeds, int64_t n_seeds,
                                 const int64_t *dev_seed_ts, int32_t *dev_rows, int64_t rows_capacity_words,
                                 rca_stats_t *stats) {
    if (!e) return RCA_ERR_ARG;
    if (!pl || n_seeds < 0 || (n_seeds > 0 && !dev_seeds) || rows_capacity_words < 0 || (rows_capacity_words > 0 && !dev_rows))
        return fail(e, RCA_ERR_ARG, "rca_expand_static: null plan, seeds or row buffer");
    if (n_seeds > RCA_STATIC_MAX_SEEDS)
        return fail(e, RCA_ERR_ARG, "rca_expand_static: %lld seeds exceed the fixed scratch (RCA_STATIC_MAX_SEEDS = %d); "
                    "use rca_expand or split the seeds", (long long)n_seeds, RCA_STATIC_MAX_SEEDS);
    if (pl->with_states && n_seeds > 0 && !dev_seed_ts) return fail(e, RCA_ERR_ARG, "rca_expand_static: with_states needs seed_ts");
    if (e->device < 0 || e->device >= kMaxDevices) return fail(e, RCA_ERR_ARG, "rca_expand_static: device %d not supported", e->device);
    int rc = validate_plan(e, pl);
    if (rc) return rc;
    CK(cudaSetDevice(e->device));
    unsigned char *ss = nullptr;
    CK(static_scratch_for(e->device, &ss));
    const bool varlen = pl->kind == RCA_PLAN_VARLEN;
    const int L = varlen ? pl->max_hops : pl->n_steps;
    const int mode = !varlen ? kModeChain : (pl->node_simple ? kModeSimple : kModeRelUnique);

    // tile-state words are recognised by the call's tag, so they are not cleared between calls;
    // the array starts as zeros (static storage) and is cleared again only if the tag wraps
    unsigned epoch;
    {
        std::lock_guard<std::mutex> lock(g_static_mu);
        epoch = ++g_static_epoch[e->device];
        if ((epoch & 0x3fffffffu) == 0) {
            CK(cudaMemsetAsync(ss + kSsTstate, 0, kSsSlots * sizeof(ulonglong2), e->stream));
            epoch = g_static_epoch[e->device] = 1;
        }
    }

    // plan table and control block: built in the engine's pinned staging buffers (allocated by
    // rca_open), copied into the scratch; the control block copy is what zeroes counters and tickets
    bool arr_prune = false;
    {
        unsigned char *t = e->h_ptab;
        std::memset(t, 0, kPtBytes);
        int *ks = reinterpret_cast<int *>(t + kPtKs);
        for (int k = 0; k < kKindPad; k++) ks[k] = k <= e->n_kinds ? e->kind_start[k] : 0x7fffffff;
        if (varlen) {
            unsigned reach[RCA_MAX_KINDS];
            reach_bits(e, pl, reach);
            for (int s = 0; s <= L; s++) {
                unsigned char *row = t + kPtAct + s * RCA_MAX_KINDS;
                for (int k = 0; k < e->n_kinds; k++) {
                    const bool dead = pl->kind_excl[k] || (pl->pos_kind[s] >= 0 && pl->pos_kind[s] != k);
                    unsigned a = dead ? kActDead : 0u;
                    if (!dead && s < L && ((reach[k] >> (L - s - 1)) & 1u)) a |= kActCont;
                    if (!dead && s >= pl->min_hops && (pl->dest_kind < 0 || pl->dest_kind == k)) a |= kActEmit;
                    if (pl->kind_inter[k]) a |= kActInter;
                    row[k] = (unsigned char)a;
                }
            }
        }
        std::memcpy(t + kPtLut, e->kind_lut.data(), kLutEntries);
        arr_prune = varlen && arrival_reach(e, pl, L, t + kPtArr);
        CK(cudaMemcpyAsync(ss + kSsPtab, t, kPtBytes, cudaMemcpyHostToDevice, e->stream));
    }
    std::memset(e->h_ctl, 0, sizeof(Ctl));
    e->h_ctl->F[0] = (unsigned long long)n_seeds;
    CK(cudaEventRecord(e->ev0, e->stream));
    CK(cudaMemcpyAsync(ss + kSsCtl, e->h_ctl, sizeof(Ctl), cudaMemcpyHostToDevice, e->stream));
    CK(cudaEventRecord(e->ev_lv[0], e->stream));
    int launches = 0;
    for (int s = 1; s <= L; s++) {
        LevelParams p;
        std::memset(&p, 0, sizeof p);
        int dir;
        if (varlen) {
            dir = pl->direction;
            p.rng = dir == RCA_DIR_OUT ? e->out_rng : dir == RCA_DIR_IN ? e->in_rng : e->und_rng;
            p.entity_only = pl->entity_only;
        } else {
            const int t = pl->step_etype[s - 1];
            dir = pl->step_dir[s - 1];
            const int rowk = dir == RCA_DIR_OUT ? e->et_src[t] : e->et_dst[t];
            p.typed = 1;
            p.rng = (dir == RCA_DIR_OUT ? e->typed_out + e->typed_out_base[t] : e->typed_in + e->typed_in_base[t]);
            p.row_lo = e->kind_start[rowk];
            p.row_hi = e->kind_start[rowk + 1];
        }
        p.adj1 = e->out_dst;
        p.adj2 = dir == RCA_DIR_IN ? e->in_adj : e->und_adj;
        p.n_nodes = e->n_nodes;
        p.s = s;
        p.chain_flag = s < L ? 1 : 2;
        p.need_inter = varlen && pl->has_inter;
        p.with_states = pl->with_states;
        p.count_only = pl->count_only;
        p.row_len = row_len_of(pl, s);
        for (int j = 0; j <= RCA_MAX_HOPS; j++) p.row_len_lv[j] = row_len_of(pl, j);
        p.epoch = epoch;
        p.lut_shift = e->lut_shift;
        p.seeds = dev_seeds; p.seed_ts = (const long long *)dev_seed_ts;
        p.frontier = reinterpret_cast<int4 *>(ss + kSsFrontier); p.fcap = RCA_STATIC_MAX_FRONTIER;
        p.out = dev_rows; p.ocap = (unsigned long long)rows_capacity_words;
        p.ctl = reinterpret_cast<Ctl *>(ss + kSsCtl);
        p.tstate = reinterpret_cast<ulonglong2 *>(ss + kSsTstate);
        p.ptab = ss + kSsPtab;
        p.rel_et = dir == RCA_DIR_OUT ? e->rel_et : dir == RCA_DIR_IN ? e->in_et : e->und_et;
        p.arr_prune = arr_prune;
        p.st_head = e->st_head; p.n_head = e->n_head; p.st_off = e->st_off; p.st_rec = e->st_rec;
        const KernelSet ks = dir == RCA_DIR_OUT ? pick_kernels<false>(mode, s - 1) : pick_kernels<true>(mode, s - 1);
        const size_t dyn = (size_t)kWarps * warp_smem_ints(s) * sizeof(int);
        int grid = max_grid(e, ks.fused, dyn);
        if (s == 1) grid = (int)std::min<long long>(grid, std::max<long long>(1, (n_seeds + kTile * kWarps - 1) / (kTile * kWarps)));
        ks.fused<<<grid, kThreads, dyn, e->stream>>>(p);
        launches++;
        CK(cudaGetLastError());
        CK(cudaEventRecord(e->ev_lv[s], e->stream));
    }
    CK(cudaEventRecord(e->ev1, e->stream));
    CK(cudaMemcpyAsync(e->h_ctl, ss + kSsCtl, sizeof(Ctl), cudaMemcpyDeviceToHost, e->stream));
    CK(cudaStreamSynchronize(e->stream));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e->ev0, e->ev1));
    const unsigned ov = e->h_ctl->overflow;
    if (ov & 1u)
        return fail(e, RCA_ERR_NOMEM, "rca_expand_static: the plan needs more than the %d partial paths the fixed scratch holds "
                    "(RCA_STATIC_MAX_FRONTIER); use rca_expand or split the seeds", RCA_STATIC_MAX_FRONTIER);
    if (ov & 2u)
        return fail(e, RCA_ERR_NOMEM, "rca_expand_static: the row buffer (%lld words) is too small for the result",
                    (long long)rows_capacity_words);
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        stats->frontier[0] = n_seeds;
        for (int s = 1; s <= L; s++) {
            stats->frontier[s] = (int64_t)e->h_ctl->F[s];
            stats->scanned[s] = (int64_t)e->h_ctl->E[s];
            stats->emitted[s] = (int64_t)e->h_ctl->N[s];
            stats->total += stats->emitted[s];
            float t = 0.f;
            if (cudaEventElapsedTime(&t, e->ev_lv[s - 1], e->ev_lv[s]) == cudaSuccess) stats->level_ms[s] = t;
        }
        for (int s = 0; s <= L; s++) stats->row_len[s] = row_len_of(pl, s);
        stats->kernel_ms = ms;
        stats->launches = launches;
    }
    return RCA_OK;
}
