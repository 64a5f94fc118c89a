// rca_kernels.cuh - sm_100a kernels of the path engine.
//
// One kernel does all the graph work: expand_level<> turns the frontier of partial paths at depth
// s-1 into the frontier at depth s and, in the same pass, emits the finished path instances of
// length s.  It replaces what Neo4j does for the reference's MATCH statements
// (find_metapath/find_srckind_metapath_neo4j.py:95-129, generate_query/generate_query.py:241-260)
// and, with `with_states`, the per-node state query of check_state/analyze_root_cause.py:70-79.
//
// Shape of one launch: persistent CTAs of 8 warps; every WARP works on its own tile of 32
// frontier entries (tiles are handed out in index order through an atomic ticket, one at a time:
// a tile claimed early but processed late would stall the look-back of every later tile), so
// there is no block-wide barrier after the set-up.
//   set-up    - one cp.async.bulk (TMA bulk copy) per CTA stages the plan's kind tables (kind
//               boundaries, per-kind attribute word, node-block -> kind table) in shared memory.
//   prologue  - lane l owns entry l of the tile: one 16 B load of the entry, one 8 B load of the
//               adjacency range of its tail node, then the parent chain is walked to recover the
//               nodes / relationships already on the path (into the warp's shared memory).
//               Warp scan of the row lengths numbers the tile's candidates 0..C-1; entries without
//               a row are compacted away.
//   pass 1    - candidates are taken 32 at a time (a "window").  The owning entry of every lane's
//               candidate comes from one shared-memory load, one warp OR-reduction and a popcount
//               (entry starts inside the window form a bit mask) - no per-candidate search.  One
//               coalesced 4 B / 8 B adjacency load per lane, then the path predicates against
//               shared memory.  Survivor ballots (continue / emit) are cached per window.
//   look-back - the tile publishes its two counts in one 128-bit word and obtains the exclusive
//               prefix over all earlier tiles (decoupled look-back, all 32 lanes polling 32
//               predecessors).  Tile order is fixed, so every survivor's slot is a pure function
//               of the input: deterministic output with no sort and no global scan kernel.  Words
//               carry the call's epoch, so the state array is never cleared between calls.
//   pass 2    - windows without survivors are skipped from the cached ballots.  Continuing
//               survivors write their 16 B frontier entry (tail, parent, rel, aux).  Emitting
//               survivors are queued in shared memory and written 32 rows at a time with every
//               lane busy: row = nodes, rels [, one HasState probe per node, missing-state mask].
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rca_b200.h"

namespace rca {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kTile = 32;             // frontier entries per tile (one warp)
constexpr int kLv = RCA_MAX_HOPS + 2;
constexpr int kLutEntries = 1024;     // node id block -> kind of its first id
constexpr int kLutWords = kLutEntries / 4;
constexpr int kKtabWords = RCA_MAX_KINDS + 4 + RCA_MAX_KINDS + kLutWords;
constexpr int kFlagWin = 64;          // windows per tile whose pass-1 ballots are cached
constexpr int kQueue = 64;            // emission queue entries per warp

// kind attribute word of a plan
constexpr unsigned kAttrExcl = 1u;
constexpr unsigned kAttrInter = 2u;
constexpr int kAttrReachShift = 8;    // bit (8 + r - 1): dest kind reachable within <= r hops

struct StRec {                        // one HasState interval: exactly one 32 B sector
    long long tmin, tmax;
    int node;
    int pad[3];
};

struct Ctl {                          // device control block, one per engine
    unsigned long long F[kLv];        // frontier size per level
    unsigned long long fbase[kLv];    // first frontier entry of the level in the arena
    unsigned long long N[kLv];        // rows emitted per level
    unsigned long long obase[kLv];    // first output word of the level in the arena
    unsigned long long E[kLv];        // candidates scanned per level
    unsigned int ticket[kLv];
    unsigned int overflow;            // bit0 frontier arena, bit1 output arena
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
    unsigned epoch;                   // tag of this call's tile-state words (never 0)
    // per-call buffers
    const int *seeds;
    const long long *seed_ts;
    int4 *frontier;
    unsigned long long fcap;
    int *out;
    unsigned long long ocap;
    Ctl *ctl;
    ulonglong2 *tstate;
    const int *ktab;                  // [nkp] kind_start (padded), [nka] attribute words, [kLutWords] block->kind
    int n_kinds, nkp, nka, lut_shift;
    // HasState intervals
    const int *st_off;
    const StRec *st_rec;
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

// 16-byte tile-state words are written and read with single vector accesses
__device__ __forceinline__ ulonglong2 ld_state(const ulonglong2 *p) {
    ulonglong2 v;
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_state(ulonglong2 *p, unsigned long long x, unsigned long long y) {
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" :: "l"(p), "l"(x), "l"(y) : "memory");
}

// kind of node v through the block table: lut[v >> sh] is the kind of the first id of v's block,
// then walk up while the next kind starts at or before v (kind_start is padded with INT_MAX)
__device__ __forceinline__ int kind_lut(const int *ks, const unsigned char *lut, int sh, int v) {
    int k = lut[v >> sh];
    while (ks[k + 1] <= v) k++;
    return k;
}

// HasState predicate for one (node, timestamp): lowest matching STATE node id (or -1) and count.
// One interval = one 32 B record, read with two 16 B loads of the same sector.
__device__ __forceinline__ void state_probe(const int *st_off, const StRec *rec, int v, long long ts, int &first,
                                            int &cnt) {
    const int b = __ldg(st_off + v), e = __ldg(st_off + v + 1);
    first = -1;
    cnt = 0;
    for (int j = b; j < e; j++) {
        const longlong2 iv = __ldg(reinterpret_cast<const longlong2 *>(rec + j));
        if (iv.x <= ts && iv.y > ts) {
            if (cnt == 0) first = __ldg(&rec[j].node);
            cnt++;
        }
    }
}

// tile-state word: x = status:2 | epoch:30 | f:32 (continuing entries), y = e (emitted rows)
constexpr unsigned kStAgg = 1u, kStPre = 2u;
__device__ __forceinline__ unsigned long long st_pack(unsigned status, unsigned epoch, unsigned long long f) {
    return ((unsigned long long)status << 62) | ((unsigned long long)(epoch & 0x3fffffffu) << 32) | (f & 0xffffffffull);
}

// Decoupled look-back over the tiles of one level; called by a whole warp for its own tile.
__device__ __forceinline__ void tile_lookback(ulonglong2 *ts, unsigned tile, unsigned epoch, unsigned long long agg_f,
                                              unsigned long long agg_e, unsigned long long &pre_f,
                                              unsigned long long &pre_e) {
    const unsigned lane = threadIdx.x & 31;
    if (tile == 0) {
        if (lane == 0) st_state(ts, st_pack(kStPre, epoch, agg_f), agg_e);
        pre_f = pre_e = 0;
        return;
    }
    if (lane == 0) st_state(ts + tile, st_pack(kStAgg, epoch, agg_f), agg_e);
    unsigned long long run_f = 0, run_e = 0;
    int base = (int)tile - 1;
    while (true) {
        const int t = base - (int)lane;
        unsigned status = kStPre;                           // virtual zero prefix before tile 0
        unsigned long long f = 0, e = 0;
        if (t >= 0) {
            ulonglong2 w;
            do {
                w = ld_state(ts + t);
                status = (unsigned)(w.x >> 62);
            } while (status == 0 || ((unsigned)(w.x >> 32) & 0x3fffffffu) != (epoch & 0x3fffffffu));
            f = w.x & 0xffffffffull;
            e = w.y;
        }
        const unsigned m = __ballot_sync(0xffffffffu, status == kStPre);
        const int first = m ? (__ffs(m) - 1) : 32;
        if ((int)lane > first) { f = 0; e = 0; }
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
    if (lane == 0) st_state(ts + tile, st_pack(kStPre, epoch, run_f + agg_f), run_e + agg_e);
}

// ------------------------------------------------------------------------------------------------
// Per-warp shared memory (carved from dynamic shared memory): fixed part, then 2*s rows of 32 ints.
struct WarpSm {
    int coff[kTile + 34];             // first candidate of each compacted entry, then C, then INT_MAX
    int crow[kTile];                  // adjacency slot of a candidate = crow[o] + c
    int cidx[kTile];                  // lane (entry index within the tile) of compacted entry o
    int aux[kTile];
    int root[kTile];                  // seed index of the path
    unsigned flag[3][kFlagWin];       // per window: continue bits, emit bits, owner before the window
    int q_o[kQueue], q_v[kQueue], q_r[kQueue];
};
__host__ __device__ constexpr int warp_smem_ints(int s) { return (int)(sizeof(WarpSm) / 4) + 2 * s * kTile; }

// Everything one candidate needs to know about the plan at this level.
struct EvalCtx {
    const int *ks;
    const unsigned *kattr;
    const unsigned char *lut;
    const int *anc_n, *anc_r;         // [s][kTile] each, in the warp's shared memory
    const int *aux;
    int lut_shift, s, q, min_hops, max_hops, node_simple, dest_kind, has_inter, pos_kind_cur;
};

// flags: bit0 continue, bit1 emit.  o = compacted owner entry, (v, r) = neighbour and relationship.
template <bool VARLEN>
__device__ __forceinline__ int eval_candidate(const EvalCtx &x, int o, int v, int r, int &naux) {
    naux = 0;
    if (!VARLEN) return (x.s < x.max_hops) ? 1 : 2;
    bool dead = false;
    if (x.node_simple) {
        for (int d = 0; d <= x.q; d++) dead |= (x.anc_n[d * kTile + o] == v);
    } else {
        for (int d = 1; d <= x.q; d++) dead |= (x.anc_r[d * kTile + o] == r);
    }
    if (dead) return 0;
    const int k = kind_lut(x.ks, x.lut, x.lut_shift, v);
    const unsigned at = x.kattr[k];
    if ((at & kAttrExcl) || (x.pos_kind_cur >= 0 && x.pos_kind_cur != k)) return 0;
    const int inter_so_far = x.aux[o] >> 16;
    naux = k | ((inter_so_far | ((at & kAttrInter) ? 1 : 0)) << 16);
    int fl = 0;
    if (x.s >= x.min_hops && (x.dest_kind < 0 || k == x.dest_kind) && (!x.has_inter || inter_so_far)) fl |= 2;
    if (x.s < x.max_hops && ((at >> (kAttrReachShift + x.max_hops - x.s - 1)) & 1u)) fl |= 1;
    return fl;
}

// ADJ2: adjacency entries are (neighbour, rel) pairs (in / undirected tables); otherwise 4-byte
// out_dst entries whose slot is the relationship id.   VARLEN: variable-length plan predicates.
// Dynamic shared memory: kWarps * warp_smem_ints(s) ints.
template <bool ADJ2, bool VARLEN>
__global__ void __launch_bounds__(kThreads) expand_level(const __grid_constant__ LevelParams p) {
    extern __shared__ int s_dyn[];
    __shared__ alignas(16) int s_ktab[kKtabWords];
    __shared__ alignas(8) unsigned long long s_bar;
    __shared__ unsigned long long s_fb[kLv];

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const int s = p.s, q = s - 1;
    Ctl *ctl = p.ctl;
    // an arena overflowed at an earlier level: entries past its end were never written, the host
    // will grow the arena and run the plan again (checked before any ticket is taken)
    if (ctl->overflow) return;

    // ---- set-up: stage the plan's kind tables with one TMA bulk copy; read level bases meanwhile
    const unsigned ktab_bytes = (unsigned)(p.nkp + p.nka + kLutWords) * 4u;
    if (tid == 0) {
        mbar_init(&s_bar, 1);
        mbar_expect_tx(&s_bar, ktab_bytes);
        bulk_g2s(s_ktab, p.ktab, ktab_bytes, &s_bar);
    }
    if (tid < kLv) s_fb[tid] = ctl->fbase[tid];
    WarpSm &W = *reinterpret_cast<WarpSm *>(s_dyn + wid * warp_smem_ints(s));
    int *anc_n = reinterpret_cast<int *>(&W + 1), *anc_r = anc_n + s * kTile;
    const unsigned long long Fq = ctl->F[q];
    const unsigned ntiles = (unsigned)((Fq + kTile - 1) / kTile);
    const unsigned long long fb_s = ctl->fbase[s], ob_s = ctl->obase[s];
    // tile-state slots: levels are laid out by cumulative entry count plus one slot of slack per
    // level, so the slot ranges of different levels never overlap
    ulonglong2 *tstate = p.tstate + ((q == 0 ? 0ull : ctl->F[0] + ctl->fbase[q]) / kTile) + q;
    __syncthreads();
    mbar_wait(&s_bar, 0);

    EvalCtx x;
    x.ks = s_ktab;
    x.kattr = reinterpret_cast<const unsigned *>(s_ktab + p.nkp);
    x.lut = reinterpret_cast<const unsigned char *>(s_ktab + p.nkp + p.nka);
    x.anc_n = anc_n; x.anc_r = anc_r; x.aux = W.aux;
    x.lut_shift = p.lut_shift; x.s = s; x.q = q; x.min_hops = p.min_hops; x.max_hops = p.max_hops;
    x.node_simple = p.node_simple; x.dest_kind = p.dest_kind; x.has_inter = p.has_inter;
    x.pos_kind_cur = p.pos_kind_cur;

    unsigned long long scanned = 0;
    unsigned ov = 0;
    while (true) {
        unsigned tile = 0;
        if (lane == 0) tile = atomicAdd(&ctl->ticket[s], 1u);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= ntiles) break;

        // ---------------------------------------------------------------- prologue: row range per entry
        const unsigned long long i = (unsigned long long)tile * kTile + lane;
        int deg = 0, beg = 0, tail = 0, par = 0, rel = 0, aux = 0;
        if (i < Fq) {
            bool live = true;
            if (q == 0) {
                tail = __ldg(p.seeds + i);
                par = (int)i;
                live = tail >= 0 && tail < p.n_nodes;
                if (live && VARLEN) {
                    const int k = kind_lut(x.ks, x.lut, x.lut_shift, tail);
                    aux = k;
                    live = !(x.kattr[k] & kAttrExcl) && (p.pos_kind_prev < 0 || p.pos_kind_prev == k);
                }
            } else {
                const int4 e = p.frontier[s_fb[q] + i];
                tail = e.x; par = e.y; rel = e.z; aux = e.w;
            }
            if (live) {
                if (p.typed) {
                    if (tail >= p.row_lo && tail < p.row_hi) {
                        const int2 rg = __ldg(p.rng + (tail - p.row_lo));
                        beg = rg.x; deg = rg.y - rg.x;
                    }
                } else {
                    const int2 rg = __ldg(p.rng + tail);
                    beg = rg.x;
                    deg = (p.entity_only ? rg.y : __ldg(p.rng + tail + 1).x) - rg.x;
                }
            }
        }
        // ---- warp scan of the row lengths; entries without a row are compacted away
        int incl = deg;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        const int C = __shfl_sync(0xffffffffu, incl, 31);
        const unsigned nzmask = __ballot_sync(0xffffffffu, deg > 0);
        const int nz = __popc(nzmask);
        __syncwarp();                                 // the previous tile's readers are done
        if (deg > 0) {
            const int o = __popc(nzmask & lt), off = incl - deg;
            W.coff[o] = off;
            W.crow[o] = beg - off;
            W.cidx[o] = lane;
            W.aux[o] = aux;
            anc_n[q * kTile + o] = tail;              // nodes / relationships already on this path
            if (q > 0) {
                anc_r[q * kTile + o] = rel;
                for (int d = q - 1; d >= 1; d--) {
                    const int4 a = p.frontier[s_fb[d] + (unsigned)par];
                    anc_n[d * kTile + o] = a.x;
                    anc_r[d * kTile + o] = a.z;
                    par = a.y;
                }
                anc_n[o] = __ldg(p.seeds + par);
            }
            W.root[o] = par;
        }
        W.coff[nz + lane] = (lane == 0) ? C : 0x7fffffff;       // coff[nz] = C, then 31 sentinels
        if (lane < 2) W.coff[nz + 32 + lane] = 0x7fffffff;
        __syncwarp();
        scanned += (unsigned long long)C;

        // owner of every lane's candidate in window `it`, given the owner jlo of the candidate
        // just before the window: count the entry starts at or before the lane's candidate
        auto owners = [&](int it, int &jlo) -> int {
            const unsigned d = (unsigned)(W.coff[jlo + 1 + lane] - (it << 5));
            const unsigned mask = __reduce_or_sync(0xffffffffu, d < 32u ? (1u << d) : 0u);
            const int o = jlo + __popc(mask & (0xffffffffu >> (31 - lane)));
            jlo += __popc(mask);
            return o;
        };
        auto fetch = [&](int o, int c, int &v, int &r) {
            const int slot = W.crow[o] + c;
            if (ADJ2) {
                const int2 a = __ldg(p.adj2 + slot);
                v = a.x; r = a.y;
            } else {
                v = __ldg(p.adj1 + slot);
                r = slot;
            }
        };

        // ---------------------------------------------------------------- pass 1: classify and count
        const int n_it = (C + 31) >> 5;
        int cf = 0, ce = 0, jcap = 0;
        {
            int jlo = 0;
            for (int it = 0; it < n_it; ++it) {
                if (it == kFlagWin) jcap = jlo;
                const int jstart = jlo;
                const int o = owners(it, jlo);
                const int c = (it << 5) + lane;
                int fl = 0;
                if (c < C) {
                    int v, r, na;
                    fetch(o, c, v, r);
                    fl = eval_candidate<VARLEN>(x, o, v, r, na);
                }
                const unsigned bf = __ballot_sync(0xffffffffu, fl & 1), be = __ballot_sync(0xffffffffu, fl & 2);
                cf += __popc(bf);
                ce += __popc(be);
                if (it < kFlagWin && lane == 0) { W.flag[0][it] = bf; W.flag[1][it] = be; W.flag[2][it] = (unsigned)jstart; }
            }
        }
        unsigned long long run_f, run_e;
        tile_lookback(tstate, tile, p.epoch, (unsigned long long)cf, (unsigned long long)ce, run_f, run_e);
        if (run_f + cf > 0xffffffffull) ov |= 1u;     // the 32-bit count field of the state word
        if (fb_s + run_f + cf > p.fcap) ov |= 1u;
        if (ob_s + (run_e + ce) * (unsigned long long)p.row_len > p.ocap) ov |= 2u;
        if (tile == ntiles - 1 && lane == 0) {        // the last tile knows the level totals
            ctl->F[s] = run_f + cf;
            ctl->N[s] = run_e + ce;
            ctl->fbase[s + 1] = fb_s + run_f + cf;
            ctl->obase[s + 1] = ob_s + (run_e + ce) * (unsigned long long)p.row_len;
        }
        __syncwarp();

        // ---------------------------------------------------------------- pass 2: write survivors
        // rows are queued and written 32 at a time so that every lane has a row to build
        int qn = 0;
        auto flush = [&](int n) {                     // emit queue entries 0..n-1 as rows run_e..run_e+n-1
            if (lane < n) {
                const int o = W.q_o[lane], v = W.q_v[lane], r = W.q_r[lane];
                const unsigned long long w0 = ob_s + (run_e + lane) * (unsigned long long)p.row_len;
                if (w0 + p.row_len <= p.ocap) {
                    int *row = p.out + w0;
                    for (int d = 0; d <= q; d++) row[d] = anc_n[d * kTile + o];
                    row[s] = v;
                    for (int d = 1; d <= q; d++) row[s + d] = anc_r[d * kTile + o];
                    row[2 * s] = r;
                    if (p.with_states) {
                        const long long ts = __ldg(p.seed_ts + W.root[o]);
                        unsigned miss = 0;
                        for (int d = 0; d <= s; d++) {
                            int first, cnt;
                            const int nd = (d <= q) ? anc_n[d * kTile + o] : v;
                            state_probe(p.st_off, p.st_rec, nd, ts, first, cnt);
                            row[2 * s + 1 + d] = first;
                            miss |= (cnt == 0 ? 1u : 0u) << d;
                        }
                        row[3 * s + 2] = (int)miss;
                    }
                }
            }
            run_e += n;
        };
        {
            int jlo = jcap;
            for (int it = 0; it < n_it; ++it) {
                unsigned bf, be;
                int o = 0, v = 0, r = 0, na = 0;
                const int c = (it << 5) + lane;
                if (it < kFlagWin) {
                    bf = W.flag[0][it]; be = W.flag[1][it];
                    if (!(bf | be)) continue;
                    int j = (int)W.flag[2][it];
                    o = owners(it, j);
                    if (((bf | be) >> lane) & 1u) {
                        fetch(o, c, v, r);
                        if (VARLEN && ((bf >> lane) & 1u)) {     // a continuing entry carries its kind and interior flag
                            const int k = kind_lut(x.ks, x.lut, x.lut_shift, v);
                            na = k | (((W.aux[o] >> 16) | ((x.kattr[k] & kAttrInter) ? 1 : 0)) << 16);
                        }
                    }
                } else {                                          // beyond the ballot cache: classify again
                    o = owners(it, jlo);
                    int fl = 0;
                    if (c < C) {
                        fetch(o, c, v, r);
                        fl = eval_candidate<VARLEN>(x, o, v, r, na);
                    }
                    bf = __ballot_sync(0xffffffffu, fl & 1); be = __ballot_sync(0xffffffffu, fl & 2);
                }
                if ((bf >> lane) & 1u) {
                    const unsigned long long pos = fb_s + run_f + __popc(bf & lt);
                    if (pos < p.fcap) p.frontier[pos] = make_int4(v, (int)(tile * kTile + W.cidx[o]), r, na);
                }
                run_f += __popc(bf);
                if (be) {
                    if ((be >> lane) & 1u) {
                        const int k = qn + __popc(be & lt);
                        W.q_o[k] = o; W.q_v[k] = v; W.q_r[k] = r;
                    }
                    qn += __popc(be);
                    __syncwarp();
                    if (qn >= 32) {
                        flush(32);
                        __syncwarp();
                        const int rest = qn - 32;
                        int to = 0, tv = 0, tr = 0;
                        if (lane < rest) { to = W.q_o[32 + lane]; tv = W.q_v[32 + lane]; tr = W.q_r[32 + lane]; }
                        __syncwarp();
                        if (lane < rest) { W.q_o[lane] = to; W.q_v[lane] = tv; W.q_r[lane] = tr; }
                        qn = rest;
                        __syncwarp();
                    }
                }
            }
            if (qn) flush(qn);
        }
    }
    if (lane == 0) {
        if (scanned) atomicAdd(&ctl->E[s], scanned);
        if (ov) atomicOr(&ctl->overflow, ov);
    }
}

// Stand-alone HasState predicate (analyze_root_cause.py:70-79), one thread per (entity, timestamp).
__global__ void state_lookup(const int *st_off, const StRec *rec, int n_nodes, const int *nodes, const long long *ts,
                             long long n, int *out_first, int *out_cnt, int *out_ids) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int v = nodes[i], c = 0, f = -1;
    if (v >= 0 && v < n_nodes) {
        const long long t = ts[i];
        for (int j = st_off[v]; j < st_off[v + 1]; j++) {
            if (rec[j].tmin <= t && rec[j].tmax > t) {
                const int sn = rec[j].node;
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
