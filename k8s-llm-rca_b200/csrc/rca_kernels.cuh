// rca_kernels.cuh - sm_100a kernels of the path engine.
//
// One kernel does all the graph work: expand_level<> turns the frontier of partial paths at depth
// s-1 into the frontier at depth s and, in the same pass, emits the finished path instances of
// length s.  It replaces what Neo4j does for the reference's MATCH statements
// (find_metapath/find_srckind_metapath_neo4j.py:95-129, generate_query/generate_query.py:241-260)
// and, with `with_states`, the per-node state query of check_state/analyze_root_cause.py:70-79.
//
// Shape of one launch (persistent CTAs, 256 threads, tiles of 256 frontier entries handed out in
// index order through an atomic ticket):
//   prologue  - each thread owns one frontier entry: it walks the parent chain to recover the
//               nodes / relationships already on the path (into shared memory) and fetches the
//               adjacency range (begin, end) of the entry's tail node.  Kind boundaries and the
//               per-kind attribute word of the plan were staged into shared memory by one
//               cp.async.bulk (TMA bulk copy) per CTA.
//   scan      - block-wide exclusive scan of the row lengths: the tile's candidates (one per
//               adjacency entry) are numbered 0..C-1, so hub rows spread over all 256 threads.
//   pass 1    - every thread evaluates candidates c, c+256, ... : binary search for the owning
//               entry, one coalesced 4 B / 8 B adjacency load, then the path predicates against
//               shared memory; counts survivors (continue / emit).
//   look-back - the tile publishes its two counts in one 64-bit word and obtains the exclusive
//               prefix over all earlier tiles (decoupled look-back); tile order is fixed, so
//               every survivor's slot is a pure function of the input - deterministic output
//               with no sort and no global scan kernel.
//   pass 2    - candidates are evaluated again (rows are L1/L2-hot), ranked with warp ballots,
//               and written: a 16 B frontier entry (tail, parent, rel, aux) and/or the finished
//               row (nodes, rels[, state ids, missing-state mask]).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rca_b200.h"

namespace rca {

constexpr int kThreads = 256;
constexpr int kTile = 256;            // frontier entries per tile (== threads)
constexpr int kLv = RCA_MAX_HOPS + 2;

// kind attribute word of a plan
constexpr unsigned kAttrExcl = 1u;
constexpr unsigned kAttrInter = 2u;
constexpr int kAttrReachShift = 8;    // bit (8 + r - 1): dest kind reachable within <= r hops

struct Ctl {                          // device control block, one per engine
    unsigned long long F[kLv];        // frontier size per level
    unsigned long long fbase[kLv];    // first frontier entry of the level in the arena
    unsigned long long N[kLv];        // rows emitted per level
    unsigned long long obase[kLv];    // first output word of the level in the arena
    unsigned long long E[kLv];        // candidates scanned per level
    unsigned int ticket[kLv];
    unsigned int overflow;            // bit0 frontier arena, bit1 output arena, bit2 > 2^31-1 per level
    unsigned int pad;
};

struct LevelParams {
    // adjacency used at this level
    const int2 *rng;                  // any: [V+1] (begin, entity_end); typed: (begin, end) per node of row kind
    const int *adj1;                  // out_dst (rel id == slot)
    const int2 *adj2;                 // in_adj / und_adj: (neighbour, rel)
    int typed;                        // rng is a typed table
    int row_lo, row_hi;               // typed: node-id range of the row kind
    int entity_only;
    int n_nodes;
    // plan
    int s;                            // level being produced (1..)
    int varlen;
    int min_hops, max_hops;           // chain: both == n_steps
    int node_simple, dest_kind, has_inter;
    int pos_kind_prev, pos_kind_cur;  // required kind at positions s-1 (seeds only) and s; -1 any
    int with_states;
    int row_len;                      // words per emitted row at this level
    // per-call buffers
    const int *seeds;
    const long long *seed_ts;
    int4 *frontier;
    unsigned long long fcap;
    int *out;
    unsigned long long ocap;
    Ctl *ctl;
    unsigned long long *tstate;
    const int *ktab;                  // [nkp] kind_start (padded) then [nka] attribute words
    int n_kinds, nkp, nka;
    // HasState intervals
    const int *st_off;
    const long long *st_tmin, *st_tmax;
    const int *st_node;
};

// ------------------------------------------------------------------------------------------------
// small device helpers

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

// TMA bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP).  bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned phase) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" :: "r"(smem_u32(bar)), "r"(phase) : "memory");
}

__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// kind of node v: largest k with kind_start[k] <= v (kind_start in shared memory)
__device__ __forceinline__ int kind_of(const int *ks, int nk, int v) {
    int lo = 0, hi = nk;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (ks[mid] <= v) lo = mid; else hi = mid;
    }
    return lo;
}

// HasState predicate for one (node, timestamp): lowest matching STATE node id (or -1) and count
__device__ __forceinline__ void state_probe(const int *st_off, const long long *tmin, const long long *tmax,
                                            const int *st_node, int v, long long ts, int &first, int &cnt) {
    int b = __ldg(st_off + v), e = __ldg(st_off + v + 1);
    first = -1;
    cnt = 0;
    for (int j = b; j < e; j++) {
        if (__ldg(tmin + j) <= ts && __ldg(tmax + j) > ts) {
            if (cnt == 0) first = __ldg(st_node + j);
            cnt++;
        }
    }
}

constexpr unsigned long long kStAgg = 1ull << 62, kStPre = 2ull << 62, kStMask = 3ull << 62;
constexpr unsigned long long kHalf = (1ull << 31) - 1;

// Decoupled look-back over tiles of one level; called by warp 0 with all 32 lanes.  Counts f
// (continuing entries) and e (emitted rows) travel in one word: status:2 | f:31 | e:31.
__device__ __forceinline__ void tile_lookback(unsigned long long *ts, unsigned tile, unsigned long long agg_f,
                                              unsigned long long agg_e, unsigned long long &pre_f,
                                              unsigned long long &pre_e) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned long long mine = (agg_f << 31) | agg_e;
    if (tile == 0) {
        if (lane == 0) st_relaxed(ts, kStPre | mine);
        pre_f = pre_e = 0;
        return;
    }
    if (lane == 0) st_relaxed(ts + tile, kStAgg | mine);
    unsigned long long run_f = 0, run_e = 0;
    int base = (int)tile - 1;
    while (true) {
        int t = base - (int)lane;
        unsigned long long w = kStPre;                      // virtual zero prefix before tile 0
        if (t >= 0) {
            do { w = ld_relaxed(ts + t); } while ((w & kStMask) == 0);
        }
        unsigned m = __ballot_sync(0xffffffffu, (w & kStMask) == kStPre);
        int first = m ? (__ffs(m) - 1) : 32;
        unsigned long long f = ((int)lane <= first) ? ((w >> 31) & kHalf) : 0ull;
        unsigned long long e = ((int)lane <= first) ? (w & kHalf) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            f += __shfl_xor_sync(0xffffffffu, f, o);
            e += __shfl_xor_sync(0xffffffffu, e, o);
        }
        run_f += f;
        run_e += e;
        if (m) break;
        base -= 32;
    }
    pre_f = run_f;
    pre_e = run_e;
    unsigned long long inc_f = run_f + agg_f, inc_e = run_e + agg_e;
    if (lane == 0) st_relaxed(ts + tile, kStPre | ((inc_f & kHalf) << 31) | (inc_e & kHalf));
}

// ------------------------------------------------------------------------------------------------
// ADJ2: adjacency entries are (neighbour, rel) pairs (in / undirected tables); otherwise 4-byte
// out_dst entries whose slot is the relationship id.   VARLEN: variable-length plan predicates.
template <bool ADJ2, bool VARLEN>
__global__ void __launch_bounds__(kThreads) expand_level(const __grid_constant__ LevelParams p) {
    __shared__ alignas(16) int s_ktab[RCA_MAX_KINDS + 4 + RCA_MAX_KINDS];
    __shared__ alignas(8) unsigned long long s_bar;
    __shared__ unsigned long long s_fb[kLv];
    __shared__ int s_degoff[kTile + 1];
    __shared__ int s_rowbeg[kTile];
    __shared__ int s_anc_n[RCA_MAX_HOPS][kTile];
    __shared__ int s_anc_r[RCA_MAX_HOPS][kTile];
    __shared__ int s_root[kTile];
    __shared__ int s_aux[kTile];
    __shared__ int s_wsum[2][kThreads / 32];
    __shared__ unsigned s_tile;
    __shared__ unsigned long long s_pre[2];

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int s = p.s, q = s - 1;
    Ctl *ctl = p.ctl;
    // an arena overflowed at an earlier level: entries past its end were never written, the host
    // will grow the arena and run the plan again (checked before any ticket is taken)
    if (ctl->overflow) return;

    // ---- stage the plan's kind tables with one TMA bulk copy; read level bases meanwhile
    const unsigned ktab_bytes = (unsigned)(p.nkp + p.nka) * 4u;
    if (tid == 0) {
        mbar_init(&s_bar, 1);
        mbar_expect_tx(&s_bar, ktab_bytes);
        bulk_g2s(s_ktab, p.ktab, ktab_bytes, &s_bar);
    }
    if (tid < kLv) s_fb[tid] = ctl->fbase[tid];
    const unsigned long long Fq = ctl->F[q];
    const unsigned ntiles = (unsigned)((Fq + kTile - 1) / kTile);
    const unsigned long long fb_s = ctl->fbase[s], ob_s = ctl->obase[s];
    // tile-state slots: levels are laid out by cumulative entry count, +1 slot of slack per level,
    // so the slot ranges of different levels never overlap (the array is zeroed once per call)
    unsigned long long *tstate = p.tstate + ((q == 0 ? 0ull : ctl->F[0] + ctl->fbase[q]) / kTile) + q;
    __syncthreads();
    mbar_wait(&s_bar, 0);
    const int *ks = s_ktab;
    const unsigned *kattr = reinterpret_cast<const unsigned *>(s_ktab + p.nkp);

    while (true) {
        if (tid == 0) s_tile = atomicAdd(&ctl->ticket[s], 1u);
        __syncthreads();
        const unsigned tile = s_tile;
        if (tile >= ntiles) break;

        // ---------------------------------------------------------------- prologue
        const unsigned long long i = (unsigned long long)tile * kTile + tid;
        int deg = 0, beg = 0;
        if (i < Fq) {
            int tail, par, aux = 0;
            bool live = true;
            if (q == 0) {
                tail = __ldg(p.seeds + i);
                par = (int)i;
                live = tail >= 0 && tail < p.n_nodes;
                if (live && VARLEN) {
                    int k = kind_of(ks, p.n_kinds, tail);
                    aux = k;
                    live = !(kattr[k] & kAttrExcl) && (p.pos_kind_prev < 0 || p.pos_kind_prev == k);
                }
            } else {
                int4 e = p.frontier[s_fb[q] + i];
                tail = e.x; par = e.y; aux = e.w;
                s_anc_r[q][tid] = e.z;
                for (int d = q - 1; d >= 1; d--) {
                    int4 a = p.frontier[s_fb[d] + (unsigned)par];
                    s_anc_n[d][tid] = a.x;
                    s_anc_r[d][tid] = a.z;
                    par = a.y;
                }
                s_anc_n[0][tid] = __ldg(p.seeds + par);
            }
            s_anc_n[q][tid] = tail;
            s_root[tid] = par;
            s_aux[tid] = aux;
            if (live) {
                if (p.typed) {
                    if (tail >= p.row_lo && tail < p.row_hi) {
                        int2 r = __ldg(p.rng + (tail - p.row_lo));
                        beg = r.x; deg = r.y - r.x;
                    }
                } else {
                    int2 r = __ldg(p.rng + tail);
                    beg = r.x;
                    deg = (p.entity_only ? r.y : __ldg(p.rng + tail + 1).x) - r.x;
                }
            }
        }
        s_rowbeg[tid] = beg;

        // ---------------------------------------------------------------- block exclusive scan of deg
        int incl = deg;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        if (lane == 31) s_wsum[0][wid] = incl;
        __syncthreads();
        int wbase = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; w++) wbase += (w < wid) ? s_wsum[0][w] : 0;
        s_degoff[tid] = wbase + incl - deg;
        if (tid == kThreads - 1) s_degoff[kTile] = wbase + incl;
        __syncthreads();
        const int C = s_degoff[kTile];

        // ---------------------------------------------------------------- candidate evaluation
        // returns flags: bit0 continue, bit1 emit; fills j (owning entry), v, r, naux
        auto eval = [&](int c, int &j, int &v, int &r, int &naux) -> int {
            int lo = 0, hi = kTile - 1;
            while (lo < hi) {
                int mid = (lo + hi + 1) >> 1;
                if (s_degoff[mid] <= c) lo = mid; else hi = mid - 1;
            }
            j = lo;
            const int slot = s_rowbeg[j] + (c - s_degoff[j]);
            if (ADJ2) {
                int2 a = __ldg(p.adj2 + slot);
                v = a.x; r = a.y;
            } else {
                v = __ldg(p.adj1 + slot);
                r = slot;
            }
            naux = 0;
            if (!VARLEN) return (s < p.max_hops) ? 1 : 2;
            bool dead = false;
            if (p.node_simple) {
                for (int d = 0; d <= q; d++) dead |= (s_anc_n[d][j] == v);
            } else {
                for (int d = 1; d <= q; d++) dead |= (s_anc_r[d][j] == r);
            }
            if (dead) return 0;
            const int k = kind_of(ks, p.n_kinds, v);
            const unsigned at = kattr[k];
            if ((at & kAttrExcl) || (p.pos_kind_cur >= 0 && p.pos_kind_cur != k)) return 0;
            const int paux = s_aux[j];
            const int inter_so_far = paux >> 16;
            naux = k | ((inter_so_far | ((at & kAttrInter) ? 1 : 0)) << 16);
            int fl = 0;
            if (s >= p.min_hops && (p.dest_kind < 0 || k == p.dest_kind) && (!p.has_inter || inter_so_far)) fl |= 2;
            if (s < p.max_hops && ((at >> (kAttrReachShift + p.max_hops - s - 1)) & 1u)) fl |= 1;
            return fl;
        };

        // ---------------------------------------------------------------- pass 1: count
        int cf = 0, ce = 0;
        for (int c = tid; c < C; c += kThreads) {
            int j, v, r, na;
            int fl = eval(c, j, v, r, na);
            cf += fl & 1;
            ce += fl >> 1;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cf += __shfl_xor_sync(0xffffffffu, cf, o);
            ce += __shfl_xor_sync(0xffffffffu, ce, o);
        }
        if (lane == 0) { s_wsum[0][wid] = cf; s_wsum[1][wid] = ce; }
        __syncthreads();
        if (wid == 0) {
            unsigned long long af = 0, ae = 0;
#pragma unroll
            for (int w = 0; w < kThreads / 32; w++) { af += s_wsum[0][w]; ae += s_wsum[1][w]; }
            unsigned long long pf, pe;
            tile_lookback(tstate, tile, af, ae, pf, pe);
            if (lane == 0) {
                s_pre[0] = pf; s_pre[1] = pe;
                if (C) atomicAdd(&ctl->E[s], (unsigned long long)C);
                unsigned ov = 0;
                if (pf + af > kHalf || pe + ae > kHalf) ov |= 4u;
                if (fb_s + pf + af > p.fcap) ov |= 1u;
                if (ob_s + (pe + ae) * (unsigned long long)p.row_len > p.ocap) ov |= 2u;
                if (ov) atomicOr(&ctl->overflow, ov);
                if (tile == ntiles - 1) {         // the last tile knows the level totals
                    ctl->F[s] = pf + af;
                    ctl->N[s] = pe + ae;
                    ctl->fbase[s + 1] = fb_s + pf + af;
                    ctl->obase[s + 1] = ob_s + (pe + ae) * (unsigned long long)p.row_len;
                }
            }
        }
        __syncthreads();
        unsigned long long run_f = s_pre[0], run_e = s_pre[1];

        // ---------------------------------------------------------------- pass 2: fill
        for (int cb = 0; cb < C; cb += kThreads) {
            const int c = cb + tid;
            int j = 0, v = 0, r = 0, na = 0, fl = 0;
            if (c < C) fl = eval(c, j, v, r, na);
            const unsigned bf = __ballot_sync(0xffffffffu, fl & 1), be = __ballot_sync(0xffffffffu, fl & 2);
            if (lane == 0) { s_wsum[0][wid] = __popc(bf); s_wsum[1][wid] = __popc(be); }
            __syncthreads();
            int wf = 0, we = 0, tf = 0, te = 0;
#pragma unroll
            for (int w = 0; w < kThreads / 32; w++) {
                int a = s_wsum[0][w], b = s_wsum[1][w];
                wf += (w < wid) ? a : 0; we += (w < wid) ? b : 0;
                tf += a; te += b;
            }
            const unsigned lt = (1u << lane) - 1u;
            if (fl & 1) {
                unsigned long long pos = fb_s + run_f + wf + __popc(bf & lt);
                if (pos < p.fcap) p.frontier[pos] = make_int4(v, (int)(tile * kTile + j), r, na);
            }
            if (fl & 2) {
                unsigned long long w0 = ob_s + (run_e + we + __popc(be & lt)) * (unsigned long long)p.row_len;
                if (w0 + p.row_len <= p.ocap) {
                    int *row = p.out + w0;
                    for (int d = 0; d <= q; d++) row[d] = s_anc_n[d][j];
                    row[s] = v;
                    for (int d = 1; d <= q; d++) row[s + d] = s_anc_r[d][j];
                    row[2 * s] = r;
                    if (p.with_states) {
                        const long long ts = __ldg(p.seed_ts + s_root[j]);
                        unsigned miss = 0;
                        for (int d = 0; d <= s; d++) {
                            int first, cnt;
                            const int nd = (d <= q) ? s_anc_n[d][j] : v;
                            state_probe(p.st_off, p.st_tmin, p.st_tmax, p.st_node, nd, ts, first, cnt);
                            row[2 * s + 1 + d] = first;
                            miss |= (cnt == 0 ? 1u : 0u) << d;
                        }
                        row[3 * s + 2] = (int)miss;
                    }
                }
            }
            run_f += tf;
            run_e += te;
            __syncthreads();
        }
    }
}

// Stand-alone HasState predicate (analyze_root_cause.py:70-79), one thread per (entity, timestamp).
__global__ void state_lookup(const int *st_off, const long long *tmin, const long long *tmax, const int *st_node,
                             int n_nodes, const int *nodes, const long long *ts, long long n, int *out_first,
                             int *out_cnt, int *out_ids) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int v = nodes[i], c = 0, f = -1;
    if (v >= 0 && v < n_nodes) {
        const long long t = ts[i];
        for (int j = st_off[v]; j < st_off[v + 1]; j++) {
            if (tmin[j] <= t && tmax[j] > t) {
                int sn = st_node[j];
                if (c == 0) f = sn;
                if (out_ids && c < RCA_STATE_LIMIT) out_ids[i * RCA_STATE_LIMIT + c] = sn;
                c++;
            }
        }
    }
    if (out_ids) for (int k = c < RCA_STATE_LIMIT ? c : RCA_STATE_LIMIT; k < RCA_STATE_LIMIT; k++) out_ids[i * RCA_STATE_LIMIT + k] = -1;
    out_first[i] = f;
    out_cnt[i] = c;
}

}  // namespace rca
