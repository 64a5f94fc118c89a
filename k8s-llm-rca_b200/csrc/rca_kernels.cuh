// rca_kernels.cuh - sm_100a kernels of the path engine.
//
// One step of a plan ("level s") turns the frontier of partial paths at depth s-1 into the
// frontier at depth s and emits the finished path instances of length s.  It replaces what Neo4j
// does for the reference's MATCH statements (find_metapath/find_srckind_metapath_neo4j.py:95-129,
// generate_query/generate_query.py:241-260) and, with `with_states`, the per-node state query of
// check_state/analyze_root_cause.py:70-79.
//
// Work decomposition (all variants): every WARP owns a tile of 32 frontier entries.
//   prologue  - lane l owns entry l: one 16 B load of the entry, one 8 B load of the adjacency
//               range of its tail node, then the parent chain is walked to recover the nodes /
//               relationships already on the path (into the warp's shared memory).  A warp scan of
//               the row lengths numbers the tile's candidates 0..C-1 (so hub rows spread over all
//               lanes); entries without a row are compacted away.
//   windows   - candidates are taken 32 at a time.  The owning entry of every lane's candidate
//               comes from one shared-memory load, one warp OR-reduction and a popcount (entry
//               starts inside the window form a bit mask) - no per-candidate search.  Then one
//               coalesced 4 B / 8 B adjacency load per lane and the path predicates: visited test
//               against shared memory, node -> kind through a block table, and one byte of the
//               per-level kind-action table (continue / emit / interior / dead) that the host
//               derived from the plan.  Kind tables reach shared memory by cp.async.bulk (TMA).
//   slots     - survivors are ranked with ballots/popcounts inside the tile; the tile's base comes
//               from a prefix over tiles in index order, so every survivor's slot is a pure
//               function of the input: deterministic output, no sort.
//   writes    - continuing survivors write a 16 B frontier entry (tail, parent, rel, aux);
//               emitting survivors are queued in shared memory and written 32 rows at a time with
//               every lane busy: row = nodes, rels [, one HasState probe per node, missing mask].
//
// Two ways to get the tile prefix:
//   split (large frontiers): count_level -> scan_groups -> fill_level.  count_level stores per-tile
//     counts and adds them into group sums (32 tiles per group); scan_groups (one CTA) scans the
//     group sums and publishes the level totals; fill_level re-walks the tile and writes, its base
//     = group prefix + a warp reduction over the counts of the earlier tiles of its group.  No
//     inter-tile waiting anywhere.
//   fused (small seed sets, where launches dominate): fused_level does count, decoupled look-back
//     over 128-bit epoch-tagged tile states, and fill in one launch.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rca_b200.h"

namespace rca {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kTile = 32;             // frontier entries per tile (one warp)
constexpr int kGroup = 32;            // tiles per group (split path)
constexpr int kLv = RCA_MAX_HOPS + 2;
constexpr int kQueue = 64;            // emission queue entries per warp
#ifndef RCA_MINB
#define RCA_MINB 6            // resident CTAs per SM the count / fused kernels are compiled for (40 registers)
#endif
#ifndef RCA_MINB_FILL
#define RCA_MINB_FILL 5       // fill_level keeps several state-probe loads in flight: 48 registers, no spills
#endif
#ifndef RCA_MINB_UNITC
#define RCA_MINB_UNITC 6      // count_unit
#endif
#ifndef RCA_MINB_UNIT
#define RCA_MINB_UNIT 4       // fill_unit: 64 registers, every probe load of a row in flight
#endif

// plan table in device memory, staged into shared memory per CTA (all offsets multiples of 16 B)
constexpr int kKindPad = RCA_MAX_KINDS + 4;                                   // kind_start ints, INT_MAX padded
constexpr int kLutEntries = 4096;                                             // node id block -> kind of its first id
constexpr int kPtKs = 0;
constexpr int kPtAct = kPtKs + 4 * kKindPad;                                  // (RCA_MAX_HOPS+1) rows of RCA_MAX_KINDS bytes
constexpr int kPtLut = kPtAct + (RCA_MAX_HOPS + 1) * RCA_MAX_KINDS;
constexpr int kArrBytes = 2 * 256;                                            // one byte per (edge type, direction)
constexpr int kPtArr = kPtLut + kLutEntries;                                  // (RCA_MAX_HOPS+1) rows of kArrBytes
constexpr int kPtBytes = kPtArr + (RCA_MAX_HOPS + 1) * kArrBytes;
constexpr int kRelRev = 1 << 30;      // rel word of in / undirected tables: hop runs against the stored direction
constexpr int kRelMask = kRelRev - 1;

// kind-action byte of one (level, kind)
constexpr unsigned kActCont = 1u;     // a path may continue through a node of this kind at this level
constexpr unsigned kActEmit = 2u;     // a path ending at a node of this kind at this level is a result
constexpr unsigned kActInter = 4u;    // kind is one of $intermediateKinds
constexpr unsigned kActDead = 8u;     // kind may not appear at this position at all

constexpr int kModeChain = 0, kModeSimple = 1, kModeRelUnique = 2;

struct StRec {                        // one HasState interval: exactly one 32 B sector
    long long tmin, tmax;
    int node;
    int pad[3];
};

// Per entity node, 16 bytes: its first interval, the STATE node of that interval and how many more
// intervals the node has, packed so that ONE 16-byte load answers the HasState predicate for every
// node with a single interval.  (A row's probes are lane-divergent loads, and the L1 data pipe - the
// unit that limits fill_unit - spends one pass per lane and load whatever the load's width.)
//   a = tmin (44 bits) | node bits 0..19 << 44
//   b = tmax (44 bits) | node bits 20..31 << 44 | min(255, number of further intervals) << 56
// Times are milliseconds clamped to [0, 2^44-1] (1970 .. 2527); a timestamp outside [0, 2^44-2] is
// answered from the full records instead (state_probe_full), so every answer is exact.
struct StHead {
    unsigned long long a, b;          // no interval: tmin 1, tmax 0 - matches no timestamp
};
constexpr unsigned long long kT44 = (1ull << 44) - 1;
// ... and, for the nodes that have more than one interval, their SECOND and THIRD interval in the same packing, 32 bytes
// per node at an address that follows from the node id alone: a node with up to three intervals is answered by the
// header and, one round trip later, this record - where the full records need three dependent round trips (offsets,
// interval bounds, then the state node of the interval that matched).  A warp waits for its slowest lane, and among
// the 128 nodes a tile of rows probes there is always one with several intervals: the chain of fill_unit is what
// bounds it.  Nodes with four or more intervals continue in the full records from the fourth on.
struct StMore {
    unsigned long long a1, b1, a2, b2;    // no such interval: tmin 1, tmax 0
};

struct Ctl {                          // device control block, one per engine
    unsigned long long F[kLv];        // frontier size per level (level s occupies arena entries sum F[1..s-1] ..)
    unsigned long long N[kLv];        // rows emitted per level  (level s occupies output words sum N[j]*row_len[j] ..)
    unsigned long long E[kLv];        // candidates scanned per level
    unsigned long long subF, subN;    // split path: entries / rows written by earlier sub-launches of the level
    unsigned long long out_base;      // first output word of this plan (plans of one submission share the arena)
    unsigned int ticket[kLv];         // fused path: tile tickets
    unsigned int overflow;            // bit0 frontier arena, bit1 output arena
    unsigned int epoch;               // tile-state tag of this run, read when the launch parameters carry none (epoch 0):
                                      // a replayed CUDA graph has fixed parameters, the tag must change with every run
    unsigned int stop_level;          // shortest_only plans: the first level that emitted rows (0: none yet); the launches
    unsigned int pad;                 // of every later level return at once
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
    int sub;                          // a level may run as several typed sub-launches (0, 1, ..): see rca_b200.cu
    int chain_flag;                   // chain plans: 1 = every candidate continues, 2 = every candidate is a row
    int need_inter;                   // emit only if an interior node so far is of an intermediate kind
    int shortest_only;                // keep only the rows of the shortest length that has any (find_metapath's rule)
    int with_states;                  // 1: rows carry a state id per node and the missing mask, 2: the mask alone
    int count_only;                   // rows are counted, never written
    int row_len;                      // words per emitted row at this level
    int row_len_lv[RCA_MAX_HOPS + 1]; // ... and at every level (arena bases are sums over earlier levels)
    unsigned epoch;                   // fused path: tag of this call's tile-state words (never 0)
    int lut_shift;
    // per-call buffers
    const int *seeds;
    const long long *seed_ts;
    int4 *frontier;
    unsigned long long fcap;
    int *out;
    unsigned long long ocap;
    Ctl *ctl;
    ulonglong2 *tstate;               // fused path: tile states
    uint2 *tcount;                    // split path: (continue, emit) counts per tile slot
    ulonglong2 *gsum;                 // split path: sums per group of 32 tile slots (zeroed per call)
    ulonglong2 *gscan;                // split path: exclusive prefix of gsum within the level
    const unsigned char *ptab;        // plan table (layout above)
    const unsigned char *rel_et;      // edge type per SLOT of the adjacency array this level walks (null: more than 256 edge
                                      // types): the load needs the slot only, so it travels beside the adjacency load
    int arr_prune;                    // refine 'continue' by the arrival table (edge type, direction)
    // unit steps (count_unit / fill_unit): the far kind of the typed step is fixed, so its tests are per launch
    int unit_fl;                      // bit0: a surviving candidate continues, bit1: it is a row, bit2: no node of the path
                                      // before its tail can be of the far kind (the count pass skips the parent walk)
    int unit_aux;                     // kind | intermediate-kind flag << 16 of the far kind
    const int2 *unit_next;            // per node of the near kind: its one (neighbour, relationship), or (-1, -1)
    // wide frontier (see "Wide entries" below): this level's fill writes / this level's kernels read entries that
    // carry the whole path so far, the states of its nodes and the seed's timestamp
    int wide_out, wide_in;
    // count-only plans whose LAST level is one unit step: the level before it takes that step while it counts
    // (count_level only: nothing is stored, so neither its own fill nor the last level is launched)
    int la_n;
    int la_lo, la_hi;                 // node-id range of the last step's near kind
    const int2 *la_next;              // its next-hop table
    // HasState intervals
    const StHead *st_head;            // [n_head]: nodes at or above n_head have no intervals
    const StMore *st_more;            // [n_head] second and third interval (null: answer from the full records)
    int n_head;
    const int *st_off;                // [V+1]: a node's records are st_rec[st_off[v] .. st_off[v+1])
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

// Shared-memory copies of the plan tables of one launch.
struct Tables {
    const int *ks;                    // kind_start, INT_MAX padded
    const unsigned char *act_q;       // kind-action row of level s-1 (seed checks)
    const unsigned char *act_s;       // kind-action row of level s
    const unsigned char *lut;
    const unsigned char *arr;         // [2*t + d]: a path that arrived over edge type t (d = 1: against its
                                      // direction) at level s can still reach the destination kind
    int lut_shift;
};
struct TablesSm {
    alignas(16) int ks[kKindPad];
    alignas(16) unsigned char act[2 * RCA_MAX_KINDS];
    alignas(16) unsigned char lut[kLutEntries];
    alignas(16) unsigned char arr[kArrBytes];
    alignas(8) unsigned long long bar;
};

// One elected thread starts the four bulk copies; everybody waits on the mbarrier after a CTA sync.
__device__ __forceinline__ Tables stage_tables(TablesSm &sm, const LevelParams &p) {
    if (threadIdx.x == 0) {
        mbar_init(&sm.bar, 1);
        mbar_expect_tx(&sm.bar, 4 * kKindPad + 2 * RCA_MAX_KINDS + kLutEntries + kArrBytes);
        bulk_g2s(sm.ks, p.ptab + kPtKs, 4 * kKindPad, &sm.bar);
        bulk_g2s(sm.act, p.ptab + kPtAct + (p.s - 1) * RCA_MAX_KINDS, 2 * RCA_MAX_KINDS, &sm.bar);
        bulk_g2s(sm.lut, p.ptab + kPtLut, kLutEntries, &sm.bar);
        bulk_g2s(sm.arr, p.ptab + kPtArr + p.s * kArrBytes, kArrBytes, &sm.bar);
    }
    __syncthreads();
    mbar_wait(&sm.bar, 0);
    Tables t;
    t.ks = sm.ks; t.act_q = sm.act; t.act_s = sm.act + RCA_MAX_KINDS; t.lut = sm.lut; t.arr = sm.arr; t.lut_shift = p.lut_shift;
    return t;
}

// kind of node v through the block table: lut[v >> sh] is the kind of the first id of v's block,
// then walk up while the next kind starts at or before v
__device__ __forceinline__ int kind_lut(const Tables &T, int v) {
    int k = T.lut[v >> T.lut_shift];
    while (T.ks[k + 1] <= v) k++;
    return k;
}

// HasState predicate for one (node, timestamp): lowest matching STATE node id (or -1) and count.
// From the full records (every node, every timestamp):
__device__ __forceinline__ void state_probe_full(const int *st_off, const StRec *rec, int v, long long ts, int from, int &first,
                                                 int &cnt) {
    const int b = __ldg(st_off + v) + from, e = __ldg(st_off + v + 1);
    for (int j = b; j < e; j++) {
        const longlong2 x = __ldg(reinterpret_cast<const longlong2 *>(rec + j));
        if (x.x <= ts && x.y > ts) {
            if (cnt == 0) first = __ldg(&rec[j].node);
            cnt++;
        }
    }
}
// Through the packed header: ld_head issues the node's one 16 B load, head_finish consumes it (and
// walks the node's other intervals, if it has any) - callers put other loads between the two.
struct States {
    const StHead *head;
    int n_head;
    const int *off;
    const StRec *rec;
    const StMore *more = nullptr;
};
__device__ __forceinline__ void packed_interval(unsigned long long a, unsigned long long b, long long ts, int &first, int &cnt) {
    if ((long long)(a & kT44) <= ts && (long long)(b & kT44) > ts) {
        if (cnt == 0) first = (int)((a >> 44) | (((b >> 44) & 0xfffull) << 20));
        cnt++;
    }
}
__device__ __forceinline__ ulonglong2 ld_head(const States &st, int v, bool on) {
    ulonglong2 h = make_ulonglong2(1ull, 0ull);
    if (on && (unsigned)v < (unsigned)st.n_head) h = __ldg(reinterpret_cast<const ulonglong2 *>(st.head + v));
    return h;
}
__device__ __forceinline__ void head_finish(const States &st, const ulonglong2 h, int v, long long ts, int &first, int &cnt) {
    first = -1;
    cnt = 0;
    if ((unsigned long long)ts >= kT44) {             // outside the packed range: exact answer from the records
        if ((unsigned)v < (unsigned)st.n_head) state_probe_full(st.off, st.rec, v, ts, 0, first, cnt);
        return;
    }
    if ((long long)(h.x & kT44) <= ts && (long long)(h.y & kT44) > ts) {
        first = (int)((h.x >> 44) | (((h.y >> 44) & 0xfffull) << 20));
        cnt = 1;
    }
    const int further = (int)(h.y >> 56);
    if (!further) return;
    if (!st.more) { state_probe_full(st.off, st.rec, v, ts, 1, first, cnt); return; }
    unsigned long long a1, b1, a2, b2;                        // one sector, ONE 256-bit request (one pass of the L1 data pipe)
    asm("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(a1), "=l"(b1), "=l"(a2), "=l"(b2) : "l"(st.more + v));
    packed_interval(a1, b1, ts, first, cnt);
    packed_interval(a2, b2, ts, first, cnt);
    if (further > 2) state_probe_full(st.off, st.rec, v, ts, 3, first, cnt);
}
__device__ __forceinline__ void state_probe(const States &st, int v, long long ts, int &first, int &cnt) {
    head_finish(st, ld_head(st, v, true), v, ts, first, cnt);
}

// ------------------------------------------------------------------------------------------------
// Per-warp shared memory (carved from dynamic shared memory): this struct, then 2*s rows of 32 ints
// (nodes and relationships already on each path of the tile).
struct WarpSm {
    int coff[kTile + 34];             // first candidate of each compacted entry, then C, then INT_MAX
    int crow[kTile];                  // adjacency slot of a candidate = crow[o] + c
    int cidx[kTile];                  // lane (entry index within the tile) of compacted entry o
    int aux[kTile];                   // kind | interior-flag << 16 of the entry's tail
    int root[kTile];                  // seed index of the path
    int q_o[kQueue], q_v[kQueue], q_r[kQueue];
};
// ... then a staging area for 32 finished rows (3s+3 words each at most), so that rows leave the SM
// as whole 128 B lines instead of 4-byte stores scattered at row stride
__host__ __device__ constexpr int warp_smem_ints(int s, int wide = 0) {
    return (int)(sizeof(WarpSm) / 4) + 2 * s * kTile + kTile * (3 * s + 3) + 0 * wide;
}

struct Warp {                         // a warp's view of its tile
    WarpSm *W;
    int *anc_n, *anc_r, *stage;
    int lane;
    unsigned lt;
};

// Wide entries.  When the level after this one is the plan's last and a unit step (fill_unit: one lane per path), that
// level is bound by its chain of dependent loads: entry -> parent entry -> seed -> state header of the seed -> its
// further intervals.  A level whose paths have at most three nodes can hand the whole path over instead: its fill
// writes, per surviving path, two 16-byte words in two planes (each plane one coalesced stream for the reader)
//     A = (node0, node1, node2 | -1, rel1)    B = (rel2 | -1, root, aux, 0)
// so that the reader knows every node of the row after ONE coalesced read and can request the next hop, the seed's
// timestamp and all state headers at once - one round trip less, and no lane-divergent parent / seed loads.
// (A first version also carried the answers of the state predicate for the nodes already on the path - looked up
// per entry and, for the new node, per candidate in the writer's window loop; measured, the writer lost twice what
// the reader gained.)  Plane stride = the level's entry count rounded up to 32.  Nothing walks up THROUGH a wide
// level: it is always the last level stored.
__device__ __forceinline__ unsigned long long wide_stride(unsigned long long entries) { return (entries + 31ull) & ~31ull; }

// Prologue of one tile; returns the number of candidates C.
template <int MODE, bool WIDE = false>
__device__ __forceinline__ int tile_prologue(const LevelParams &p, const Tables &T, const Warp &w, unsigned tile,
                                             unsigned long long Fq, const unsigned long long *fb) {
    const int s = p.s, q = s - 1, lane = w.lane;
    WarpSm &W = *w.W;
    const unsigned long long i = (unsigned long long)tile * kTile + lane;
    int deg = 0, beg = 0, tail = 0, par = 0, rel = 0, aux = 0;
    if (i < Fq) {
        bool live = true;
        if (q == 0) {
            tail = __ldg(p.seeds + i);
            par = (int)i;
            live = tail >= 0 && tail < p.n_nodes;
            if (live && MODE != kModeChain) {
                const int k = kind_lut(T, tail);
                aux = k;
                live = !(T.act_q[k] & kActDead);
            }
        } else {
            const int4 e = p.frontier[fb[q] + i];
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
    int incl = deg;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    const int C = __shfl_sync(0xffffffffu, incl, 31);
    const unsigned nzmask = __ballot_sync(0xffffffffu, deg > 0);
    const int nz = __popc(nzmask);
    __syncwarp();                                     // the previous tile's readers are done
    if (deg > 0) {
        const int o = __popc(nzmask & w.lt), off = incl - deg;
        W.coff[o] = off;
        W.crow[o] = beg - off;
        W.cidx[o] = lane;
        W.aux[o] = aux;
        w.anc_n[q * kTile + o] = tail;
        if (q > 0) {
            w.anc_r[q * kTile + o] = rel;
            for (int d = q - 1; d >= 1; d--) {
                const int4 a = p.frontier[fb[d] + (unsigned)par];
                w.anc_n[d * kTile + o] = a.x;
                w.anc_r[d * kTile + o] = a.z;
                par = a.y;
            }
            w.anc_n[o] = __ldg(p.seeds + par);
        }
        W.root[o] = par;
    }
    W.coff[nz + lane] = (lane == 0) ? C : 0x7fffffff;            // coff[nz] = C, then sentinels
    if (lane < 2) W.coff[nz + 32 + lane] = 0x7fffffff;
    __syncwarp();
    return C;
}

// owner of every lane's candidate in window `it`, given the owner jlo of the candidate just
// before the window: count the entry starts at or before the lane's candidate
__device__ __forceinline__ int window_owner(const WarpSm &W, int it, int &jlo, int lane) {
    const unsigned d = (unsigned)(W.coff[jlo + 1 + lane] - (it << 5));
    const unsigned mask = __reduce_or_sync(0xffffffffu, d < 32u ? (1u << d) : 0u);
    const int o = jlo + __popc(mask & (0xffffffffu >> (31 - lane)));
    jlo += __popc(mask);
    return o;
}

// A lane's candidate as it comes out of the adjacency arrays.  (Fetching it one window ahead of its use
// was measured: no gain at level 1, slower at levels 2 and 3 - the loop is not waiting on this load.)
struct Cand {
    int v, rw, et;                    // neighbour, relationship word (bit 30: against the stored direction), edge type
};
template <bool ADJ2, int MODE>
__device__ __forceinline__ Cand fetch_cand(const LevelParams &p, const Warp &w, int o, int c, int C) {
    Cand x;
    x.v = 0; x.rw = 0; x.et = 0;
    if (c >= C) return x;
    const int slot = w.W->crow[o] + c;
    if (MODE != kModeChain && p.arr_prune) x.et = __ldg(p.rel_et + slot);
    if (ADJ2) {
        const int2 a = __ldg(p.adj2 + slot);
        x.v = a.x; x.rw = a.y;
    } else {
        x.v = __ldg(p.adj1 + slot);
        x.rw = slot;
    }
    return x;
}

// Classify the lane's candidate of window `it`: bit0 continue, bit1 emit.
// QT: number of earlier path nodes minus one when known at compile time (0..3), else -1.
template <bool ADJ2, int MODE, int QT>
__device__ __forceinline__ int classify(const LevelParams &p, const Tables &T, const Warp &w, int o, int c, int C,
                                        const Cand x, int &v, int &r, int &na) {
    v = 0; r = 0; na = 0;
    if (c >= C) return 0;
    v = x.v;
    r = ADJ2 ? (x.rw & kRelMask) : x.rw;
    const int against = ADJ2 ? ((x.rw >> 30) & 1) : 0, et = x.et;
    if (MODE == kModeChain) return p.chain_flag;
    bool dead = false;
    if (MODE == kModeSimple) {
        if (QT >= 0) {
#pragma unroll
            for (int d = 0; d <= QT; d++) dead |= (w.anc_n[d * kTile + o] == v);
        } else {
            for (int d = 0; d < p.s; d++) dead |= (w.anc_n[d * kTile + o] == v);
        }
    } else {
        for (int d = 1; d < p.s; d++) dead |= (w.anc_r[d * kTile + o] == r);
    }
    if (dead) return 0;
    const int k = kind_lut(T, v);
    const unsigned a = T.act_s[k];
    const int inter = w.W->aux[o] >> 16;
    na = k | ((inter | (int)((a >> 2) & 1u)) << 16);
    int fl = (int)(a & kActCont) | (((a & kActEmit) && (!p.need_inter || inter)) ? 2 : 0);
    if ((fl & 1) && p.arr_prune && !T.arr[2 * et + against]) fl &= ~1;
    return fl;
}

// Emit queue entries 0..n-1 as rows run_e .. run_e+n-1 (one row per lane; called by the whole warp).
// State probes are shared: rows written together are siblings (same seed, usually the same second
// node), so per column the lanes that hold the same node id elect one lane to probe and take its
// answer by shuffle.
template <int S>        // S: path length of the rows when known at compile time (1..4), else -1
__device__ __forceinline__ void flush_rows(const LevelParams &p, const Warp &w, int n, unsigned long long ob_s,
                                           unsigned long long run_e) {
    if (p.count_only) return;
    const int s = p.s, q = s - 1, rl = p.row_len;
    const WarpSm &W = *w.W;
    const bool active = w.lane < n;
    const unsigned long long w0 = ob_s + run_e * (unsigned long long)rl;       // first word of the n rows
    if (w0 + (unsigned long long)n * rl > p.ocap) return;        // overflow was flagged with the counts: the run is void
    int o = 0, v = 0, r = 0;
    int *row = w.stage + w.lane * rl;
    if (active) {
        o = W.q_o[w.lane]; v = W.q_v[w.lane]; r = W.q_r[w.lane];
        for (int d = 0; d <= q; d++) row[d] = w.anc_n[d * kTile + o];
        row[s] = v;
        for (int d = 1; d <= q; d++) row[s + d] = w.anc_r[d * kTile + o];
        row[2 * s] = r;
    }
    if (p.with_states) {
        const long long ts = active ? __ldg(p.seed_ts + W.root[o]) : 0;
        unsigned miss = 0;
        if constexpr (S >= 0) {
            // the kernel is latency-bound here: issue every column's offset loads, then every column's
            // record loads, back to back - two memory round trips per row instead of 2(S+1)
            const States st{p.st_head, p.n_head, p.st_off, p.st_rec, p.st_more};
            ulonglong2 h[S + 1];
#pragma unroll
            for (int d = 0; d <= S; d++) h[d] = ld_head(st, (d <= q) ? w.anc_n[d * kTile + o] : v, active);
#pragma unroll
            for (int d = 0; d <= S; d++) {
                int first, cnt;
                head_finish(st, h[d], (d <= q) ? w.anc_n[d * kTile + o] : v, ts, first, cnt);
                if (active) {
                    if (p.with_states == 1) row[2 * s + 1 + d] = first;
                    miss |= (cnt == 0 ? 1u : 0u) << d;
                }
            }
        } else {
            for (int d = 0; d <= s; d++) {
                const int nd = active ? ((d <= q) ? w.anc_n[d * kTile + o] : v) : -1;
                // same node AND same timestamp -> same answer (rows of one seed share the timestamp)
                const unsigned same = __match_any_sync(0xffffffffu, nd) & __match_any_sync(0xffffffffu, ts);
                const int leader = __ffs(same) - 1;
                int first = -1, cnt = 0;
                if (w.lane == leader && nd >= 0) state_probe(States{p.st_head, p.n_head, p.st_off, p.st_rec, p.st_more}, nd, ts, first, cnt);
                first = __shfl_sync(0xffffffffu, first, leader);
                cnt = __shfl_sync(0xffffffffu, cnt, leader);
                if (active) {
                    if (p.with_states == 1) row[2 * s + 1 + d] = first;
                    miss |= (cnt == 0 ? 1u : 0u) << d;
                }
            }
        }
        if (active) row[p.with_states == 1 ? 3 * s + 2 : 2 * s + 1] = (int)miss;        // compact rows: the mask alone
    }
    __syncwarp();
    int *dst = p.out + w0;
    for (int k = w.lane; k < n * rl; k += 32) dst[k] = w.stage[k];              // the n rows are contiguous in the output
    __syncwarp();
}

// Walk the tile's windows, classify, and write the survivors starting at the tile's bases.
template <bool ADJ2, int MODE, int QT, bool WIDE = false>
__device__ __forceinline__ void tile_fill(const LevelParams &p, const Tables &T, const Warp &w, unsigned tile, int C,
                                          unsigned long long fb_s, unsigned long long ob_s,
                                          unsigned long long run_f, unsigned long long run_e, unsigned long long wstride = 0) {
    WarpSm &W = *w.W;
    const int lane = w.lane;
    const int n_it = (C + 31) >> 5;
    int jlo = 0, qn = 0;
    for (int it = 0; it < n_it; ++it) {
        const int o = window_owner(W, it, jlo, lane);
        const Cand x = fetch_cand<ADJ2, MODE>(p, w, o, (it << 5) + lane, C);
        int v, r, na;
        const int fl = classify<ADJ2, MODE, QT>(p, T, w, o, (it << 5) + lane, C, x, v, r, na);
        const unsigned bf = __ballot_sync(0xffffffffu, fl & 1), be = __ballot_sync(0xffffffffu, fl & 2);
        if (WIDE) {
            if (fl & 1) {                             // QT + 1 = nodes on the path before this step (1 or 2)
                const unsigned long long j = run_f + __popc(bf & w.lt);
                int4 *A = p.frontier + fb_s;
                if (fb_s + 2 * wstride <= p.fcap) {   // else: flagged by the scan, the run is void
                    const int n0 = w.anc_n[o];
                    const int n1 = QT >= 1 ? w.anc_n[kTile + o] : v, n2 = QT >= 1 ? v : -1;
                    const int r1 = QT >= 1 ? w.anc_r[kTile + o] : r, r2 = QT >= 1 ? r : -1;
                    A[j] = make_int4(n0, n1, n2, r1);
                    A[wstride + j] = make_int4(r2, W.root[o], na, 0);
                }
            }
        } else if (fl & 1) {
            const unsigned long long pos = fb_s + run_f + __popc(bf & w.lt);
            if (pos < p.fcap) p.frontier[pos] = make_int4(v, (int)(tile * kTile + W.cidx[o]), r, na);
        }
        run_f += __popc(bf);
        if (be) {                                     // rows are queued and written 32 at a time
            if (fl & 2) {
                const int k = qn + __popc(be & w.lt);
                W.q_o[k] = o; W.q_v[k] = v; W.q_r[k] = r;
            }
            qn += __popc(be);
            __syncwarp();
            if (qn >= 32) {
                flush_rows<(QT >= 0 && QT <= 3) ? QT + 1 : -1>(p, w, 32, ob_s, run_e);
                run_e += 32;
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
    if (qn) flush_rows<(QT >= 0 && QT <= 3) ? QT + 1 : -1>(p, w, qn, ob_s, run_e);
}

// Walk the tile's windows and count the survivors.  With p.la_n (see LevelParams) every path that continues also
// takes the plan's last step - one load from a next-hop table - and c2 / s2 count the rows and the entries looked at.
template <bool ADJ2, int MODE, int QT>
__device__ __forceinline__ void tile_count(const LevelParams &p, const Tables &T, const Warp &w, int C, int &cf,
                                           int &ce, int &c2, int &s2) {
    const int n_it = (C + 31) >> 5;
    int jlo = 0;
    cf = 0; ce = 0; c2 = 0; s2 = 0;
    for (int it = 0; it < n_it; ++it) {
        const int o = window_owner(*w.W, it, jlo, w.lane);
        const Cand x = fetch_cand<ADJ2, MODE>(p, w, o, (it << 5) + w.lane, C);
        int v, r, na;
        const int fl = classify<ADJ2, MODE, QT>(p, T, w, o, (it << 5) + w.lane, C, x, v, r, na);
        cf += __popc(__ballot_sync(0xffffffffu, fl & 1));
        ce += __popc(__ballot_sync(0xffffffffu, fl & 2));
        if (MODE == kModeSimple && p.la_n) {
            int2 nx = make_int2(-1, -1);
            if ((fl & 1) && v >= p.la_lo && v < p.la_hi) nx = __ldg(p.la_next + (v - p.la_lo));
            bool ok = nx.x >= 0 && nx.x != v && (!p.need_inter || (na >> 16));
            if (ok) {
                if (QT >= 0) {
#pragma unroll
                    for (int d = 0; d <= QT; d++) ok &= (w.anc_n[d * kTile + o] != nx.x);
                } else {
                    for (int d = 0; d < p.s; d++) ok &= (w.anc_n[d * kTile + o] != nx.x);
                }
            }
            s2 += __popc(__ballot_sync(0xffffffffu, nx.x >= 0));
            c2 += __popc(__ballot_sync(0xffffffffu, ok));
        }
    }
}

// CTA-uniform read of the overflow flag (it can be raised while a fused launch is running, so
// every thread must act on the same value before the first CTA barrier)
__device__ __forceinline__ bool overflow_seen(const Ctl *ctl) {
    __shared__ unsigned s_ov;
    if (threadIdx.x == 0) s_ov = *reinterpret_cast<const volatile unsigned *>(&ctl->overflow);
    __syncthreads();
    return s_ov != 0;
}
// ... and of "a shorter level already has rows" (shortest_only plans: this level's launches have nothing to do)
__device__ __forceinline__ bool launch_void(const LevelParams &p) {
    __shared__ unsigned s_void;
    if (threadIdx.x == 0) {
        const volatile Ctl *c = p.ctl;
        const unsigned stop = c->stop_level;
        s_void = c->overflow != 0 || (stop != 0 && (unsigned)p.s > stop);
    }
    __syncthreads();
    return s_void != 0;
}

// Common set-up of the three level kernels: tables, per-warp shared memory, arena bases.
struct LevelCtx {
    Tables T;
    Warp w;
    unsigned long long Fq, fb_s, ob_s;
    unsigned ntiles;
    unsigned long long slot0;         // first tile slot of this level (multiple of kGroup)
};
__device__ __forceinline__ LevelCtx level_setup(const LevelParams &p, TablesSm &tsm, unsigned long long *s_fb,
                                                int *s_dyn, int wide = 0) {
    LevelCtx c;
    const int tid = threadIdx.x, s = p.s, q = s - 1;
    Ctl *ctl = p.ctl;
    if (tid < kLv) {                                  // arena base of every level below s, from the totals so far
        unsigned long long fb = 0;
        for (int j = 1; j < tid && j < s; j++) fb += ctl->F[j];
        s_fb[tid] = fb;
    }
    c.T = stage_tables(tsm, p);                       // contains the CTA barrier that also publishes s_fb
    c.w.lane = tid & 31;
    c.w.lt = (1u << c.w.lane) - 1u;
    c.w.W = reinterpret_cast<WarpSm *>(s_dyn + (tid >> 5) * warp_smem_ints(s, wide));
    c.w.anc_n = reinterpret_cast<int *>(c.w.W + 1);
    c.w.anc_r = c.w.anc_n + s * kTile;
    c.w.stage = c.w.anc_r + s * kTile;
    unsigned long long fb_s = 0, ob_s = ctl->out_base;
    for (int j = 1; j < s; j++) { fb_s += ctl->F[j]; ob_s += ctl->N[j] * (unsigned long long)p.row_len_lv[j]; }
    c.fb_s = fb_s; c.ob_s = ob_s;
    c.Fq = ctl->F[q];
    c.ntiles = (unsigned)((c.Fq + kTile - 1) / kTile);
    // tile slots: levels are laid out by cumulative entry count, rounded up to a group, plus one
    // group of slack per level, so the slot ranges of different levels never overlap
    const unsigned long long vpos = (q == 0) ? 0ull : ctl->F[0] + (fb_s - ctl->F[q]);
    c.slot0 = ((vpos / kTile + kGroup - 1) / kGroup + (unsigned long long)q) * kGroup;
    return c;
}

// ------------------------------------------------------------------------------------------------
// split path, step 1: per-tile survivor counts, added into the group sums
template <bool ADJ2, int MODE, int QT>
__global__ void __launch_bounds__(kThreads, RCA_MINB) count_level(const __grid_constant__ LevelParams p) {
    extern __shared__ int s_dyn[];
    __shared__ TablesSm tsm;
    __shared__ unsigned long long s_fb[kLv];
    if (launch_void(p)) return;
    const LevelCtx c = level_setup(p, tsm, s_fb, s_dyn);
    unsigned long long scanned = 0, rows2 = 0, scanned2 = 0;
    const unsigned nwarps = gridDim.x * kWarps;
    for (unsigned tile = blockIdx.x * kWarps + (threadIdx.x >> 5); tile < c.ntiles; tile += nwarps) {
        const int C = tile_prologue<MODE>(p, c.T, c.w, tile, c.Fq, s_fb);
        scanned += (unsigned long long)C;
        int cf, ce, c2, s2;
        tile_count<ADJ2, MODE, QT>(p, c.T, c.w, C, cf, ce, c2, s2);
        rows2 += (unsigned long long)c2;
        scanned2 += (unsigned long long)s2;
        if (c.w.lane == 0) {
            const unsigned long long slot = c.slot0 + tile;
            p.tcount[slot] = make_uint2((unsigned)cf, (unsigned)ce);
            if (cf) atomicAdd(&p.gsum[slot / kGroup].x, (unsigned long long)cf);
            if (ce) atomicAdd(&p.gsum[slot / kGroup].y, (unsigned long long)ce);
        }
    }
    if (c.w.lane == 0) {
        if (scanned) atomicAdd(&p.ctl->E[p.s], scanned);
        if (rows2) atomicAdd(&p.ctl->N[p.s + 1], rows2);      // the merged last step's rows and entries
        if (scanned2) atomicAdd(&p.ctl->E[p.s + 1], scanned2);
    }
}

// split path, step 2 (one CTA): exclusive scan of the group sums, level totals, overflow check
__global__ void __launch_bounds__(1024) scan_groups(const __grid_constant__ LevelParams p) {
    __shared__ unsigned long long s_w[2][32];
    __shared__ unsigned long long s_run[2];
    Ctl *ctl = p.ctl;
    if (ctl->overflow || (ctl->stop_level != 0 && (unsigned)p.s > ctl->stop_level)) return;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, s = p.s, q = s - 1;
    unsigned long long fb_s = 0, ob_s = ctl->out_base;
    for (int j = 1; j < s; j++) { fb_s += ctl->F[j]; ob_s += ctl->N[j] * (unsigned long long)p.row_len_lv[j]; }
    const unsigned long long Fq = ctl->F[q];
    const unsigned long long ntiles = (Fq + kTile - 1) / kTile;
    const unsigned long long vpos = (q == 0) ? 0ull : ctl->F[0] + (fb_s - ctl->F[q]);
    const unsigned long long g0 = (vpos / kTile + kGroup - 1) / kGroup + (unsigned long long)q;
    const unsigned long long ng = (ntiles + kGroup - 1) / kGroup;
    // every thread owns a contiguous run of groups: its loads are all in flight at once, then one
    // block-wide scan of the 1024 partial sums, then the thread's own prefixes
    const unsigned long long chunk = (ng + 1023) / 1024;
    const unsigned long long lo = min((unsigned long long)tid * chunk, ng), hi = min(lo + chunk, ng);
    unsigned long long a = 0, b = 0;
#pragma unroll 4
    for (unsigned long long g = lo; g < hi; g++) {
        const ulonglong2 v = p.gsum[g0 + g];
        a += v.x; b += v.y;
    }
    const unsigned long long ta = a, tb = b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long na = __shfl_up_sync(0xffffffffu, a, o), nb = __shfl_up_sync(0xffffffffu, b, o);
        if (lane >= o) { a += na; b += nb; }
    }
    if (lane == 31) { s_w[0][wid] = a; s_w[1][wid] = b; }
    __syncthreads();
    unsigned long long wa = 0, wb = 0;
    for (int k = 0; k < wid; k++) { wa += s_w[0][k]; wb += s_w[1][k]; }
    unsigned long long ea = wa + a - ta, eb = wb + b - tb;             // exclusive prefix of this thread's run
    for (unsigned long long g = lo; g < hi; g++) {
        const ulonglong2 v = p.gsum[g0 + g];
        p.gscan[g0 + g] = make_ulonglong2(ea, eb);
        ea += v.x; eb += v.y;
    }
    if (tid == 1023) { s_run[0] = ea; s_run[1] = eb; }
    __syncthreads();
    if (tid == 0) {
        // sub-launch 0 starts the level; later ones append after what the earlier ones wrote
        const unsigned long long base_f = p.sub ? ctl->F[s] : 0ull, base_n = p.sub ? ctl->N[s] : 0ull;
        ctl->subF = base_f;
        ctl->subN = base_n;
        ctl->F[s] = base_f + s_run[0];
        ctl->N[s] = base_n + s_run[1];
        unsigned ov = 0;
        const unsigned long long room = p.wide_out ? 2ull * wide_stride(base_f + s_run[0]) : base_f + s_run[0];
        if (!p.la_n && fb_s + room > p.fcap) ov |= 1u;                 // la_n: the paths are counted, never stored
        if (!p.count_only && ob_s + (base_n + s_run[1]) * (unsigned long long)p.row_len > p.ocap) ov |= 2u;
        if (ov) ctl->overflow = ov;
        if (p.shortest_only && base_n + s_run[1] > 0 && !ctl->stop_level) ctl->stop_level = (unsigned)s;
    }
}

// split path, step 3: re-walk every tile and write its survivors at (group prefix + in-group prefix)
template <bool ADJ2, int MODE, int QT, bool WIDE = false>
__global__ void __launch_bounds__(kThreads, RCA_MINB_FILL) fill_level(const __grid_constant__ LevelParams p) {
    extern __shared__ int s_dyn[];
    __shared__ TablesSm tsm;
    __shared__ unsigned long long s_fb[kLv];
    if (launch_void(p)) return;
    LevelCtx c = level_setup(p, tsm, s_fb, s_dyn, WIDE ? 1 : 0);
    c.fb_s += p.ctl->subF;                            // after the earlier sub-launches of this level
    c.ob_s += p.ctl->subN * (unsigned long long)p.row_len;
    const unsigned long long wstride = WIDE ? wide_stride(p.ctl->F[p.s]) : 0ull;     // a wide level is one launch: F[s] is final
    const unsigned nwarps = gridDim.x * kWarps;
    for (unsigned tile = blockIdx.x * kWarps + (threadIdx.x >> 5); tile < c.ntiles; tile += nwarps) {
        const unsigned long long slot = c.slot0 + tile, g = slot / kGroup;
        const unsigned in_g = (unsigned)(slot % kGroup);
        // counts of the earlier tiles of the group (slot0 is group-aligned, so they belong to this level)
        uint2 mine = make_uint2(0, 0);
        if ((unsigned)c.w.lane < in_g) mine = p.tcount[g * kGroup + c.w.lane];
        const ulonglong2 gp = p.gscan[g];
        const unsigned long long run_f = gp.x + __reduce_add_sync(0xffffffffu, mine.x);
        const unsigned long long run_e = gp.y + __reduce_add_sync(0xffffffffu, mine.y);
        const int C = tile_prologue<MODE, WIDE>(p, c.T, c.w, tile, c.Fq, s_fb);
        tile_fill<ADJ2, MODE, QT, WIDE>(p, c.T, c.w, tile, C, c.fb_s, c.ob_s, run_f, run_e, wstride);
    }
}

// ------------------------------------------------------------------------------------------------
// Unit steps.  A typed step over an edge type no node has two of on the near side (Pod -spec_nodeName->
// Node, ownerReferences, claimRef ...: the engine knows from the typed-range tables) gives every
// partial path at most ONE candidate.  Then a tile needs no scan of row lengths, no compaction, no
// windows and no emission queue - one lane, one path, from its frontier entry to its finished row -
// and, since nothing waits on a neighbour lane, every load of that lane can be in flight as early as
// its address is known: the parent walk and the state probes of the nodes already on the path run
// beside the step itself instead of after it, and the step is ONE load - for such edge types the engine
// keeps, per node of the near kind, the neighbour and the relationship themselves (LevelParams::unit_next)
// rather than a range into an adjacency array (4 dependent round trips per row, not 8).
// The kind tests are per launch, not per candidate: the far kind of the step is fixed (p.unit_fl).
// Same tile slots, same counts, same scan as count_level / fill_level; split path, levels >= 2.

struct UnitCtx {
    unsigned long long Fq, fb_s, ob_s, slot0;
    unsigned ntiles;
};
__device__ __forceinline__ UnitCtx unit_setup(const LevelParams &p, unsigned long long *s_fb) {
    UnitCtx c;
    const int tid = threadIdx.x, s = p.s, q = s - 1;
    Ctl *ctl = p.ctl;
    if (tid < kLv) {
        unsigned long long fb = 0;
        for (int j = 1; j < tid && j < s; j++) fb += ctl->F[j];
        s_fb[tid] = fb;
    }
    __syncthreads();
    unsigned long long fb_s = 0, ob_s = ctl->out_base;
    for (int j = 1; j < s; j++) { fb_s += ctl->F[j]; ob_s += ctl->N[j] * (unsigned long long)p.row_len_lv[j]; }
    c.fb_s = fb_s; c.ob_s = ob_s;
    c.Fq = ctl->F[q];
    c.ntiles = (unsigned)((c.Fq + kTile - 1) / kTile);
    const unsigned long long vpos = ctl->F[0] + (fb_s - ctl->F[q]);
    c.slot0 = ((vpos / kTile + kGroup - 1) / kGroup + (unsigned long long)q) * kGroup;
    return c;
}

// One lane's path: its nodes a[0..Q] and relationships rel[1..Q] (walked up the parent entries), the
// candidate (v, r) and what becomes of it (bit0 continue, bit1 emit).
template <int Q>
struct UnitLane {
    int a[Q + 1], rel[Q + 1];
    int v, r, root, inter, fl;
    bool has;
    // e: the lane's frontier entry (fetched one tile ahead by the callers)
    __device__ __forceinline__ void load(const LevelParams &p, const unsigned long long *fb, const int4 e, bool valid) {
        a[Q] = e.x; rel[Q] = e.z; inter = e.w >> 16;
        int2 nx = make_int2(-1, -1);
        if (valid && e.x >= p.row_lo && e.x < p.row_hi) nx = __ldg(p.unit_next + (e.x - p.row_lo));
        int par = e.y;
#pragma unroll
        for (int d = Q - 1; d >= 1; d--) {
            int4 pe = make_int4(0, 0, 0, 0);
            if (valid) pe = p.frontier[fb[d] + (unsigned)par];
            a[d] = pe.x; rel[d] = pe.z; par = pe.y;
        }
        root = par;
        a[0] = valid ? __ldg(p.seeds + par) : 0;
        rel[0] = 0;
        has = nx.x >= 0;
        v = nx.x; r = nx.y;
        bool dead = !has;
#pragma unroll
        for (int d = 0; d <= Q; d++) dead |= (a[d] == v);
        fl = dead ? 0 : ((p.unit_fl & 1) | (((p.unit_fl & 2) && (!p.need_inter || inter)) ? 2 : 0));
    }
    // the count pass when the host has shown that no node before the tail can be of the step's far kind (unit_fl bit 2):
    // the candidate can only collide with the tail itself, so the parent walk and the seed load are not needed
    __device__ __forceinline__ void load_tail_only(const LevelParams &p, const int4 e, bool valid) {
        a[Q] = e.x; inter = e.w >> 16;
        int2 nx = make_int2(-1, -1);
        if (valid && e.x >= p.row_lo && e.x < p.row_hi) nx = __ldg(p.unit_next + (e.x - p.row_lo));
        has = nx.x >= 0;
        v = nx.x; r = nx.y;
        const bool dead = !has || v == e.x;
        fl = dead ? 0 : ((p.unit_fl & 1) | (((p.unit_fl & 2) && (!p.need_inter || inter)) ? 2 : 0));
    }
    // the same from a wide entry (Q = 1 or 2): plane A holds the nodes, B (needed by the count pass only under the
    // intermediate-kind rule) the rest
    __device__ __forceinline__ void load_wide(const LevelParams &p, const int4 A, const int4 B, bool valid) {
        a[0] = A.x; a[1] = A.y;
        if (Q >= 2) a[2] = A.z;
        rel[0] = 0; rel[1] = A.w;
        if (Q >= 2) rel[2] = B.x;
        root = B.y; inter = B.z >> 16;
        int2 nx = make_int2(-1, -1);
        if (valid && a[Q] >= p.row_lo && a[Q] < p.row_hi) nx = __ldg(p.unit_next + (a[Q] - p.row_lo));
        has = nx.x >= 0;
        v = nx.x; r = nx.y;
        bool dead = !has;
#pragma unroll
        for (int d = 0; d <= Q; d++) dead |= (a[d] == v);
        fl = dead ? 0 : ((p.unit_fl & 1) | (((p.unit_fl & 2) && (!p.need_inter || inter)) ? 2 : 0));
    }
};

template <int Q, bool WIDE = false>
__global__ void __launch_bounds__(kThreads, RCA_MINB_UNITC) count_unit(const __grid_constant__ LevelParams p) {
    __shared__ unsigned long long s_fb[kLv];
    if (launch_void(p)) return;
    const UnitCtx c = unit_setup(p, s_fb);
    const int lane = threadIdx.x & 31;
    unsigned long long scanned = 0;
    const unsigned nwarps = gridDim.x * kWarps;
    const int4 *fq = p.frontier + s_fb[Q];            // wide: plane A
    const unsigned long long wstride = WIDE ? wide_stride(c.Fq) : 0ull;
    const int4 none = make_int4(0, 0, 0, 0);
    unsigned tile = blockIdx.x * kWarps + (threadIdx.x >> 5);
    int4 e_next = ((unsigned long long)tile * kTile + lane < c.Fq) ? fq[(unsigned long long)tile * kTile + lane] : none;
    for (; tile < c.ntiles; tile += nwarps) {
        const unsigned long long i = (unsigned long long)tile * kTile + lane, i2 = i + (unsigned long long)nwarps * kTile;
        const int4 e = e_next;
        e_next = i2 < c.Fq ? fq[i2] : none;           // the next tile's entries travel while this tile is worked on
        UnitLane<Q> u;
        if constexpr (WIDE) {
            int4 B = none;
            if (p.need_inter && i < c.Fq) B = fq[wstride + i];
            u.load_wide(p, e, B, i < c.Fq);
        } else if (p.unit_fl & 4) {
            u.load_tail_only(p, e, i < c.Fq);
        } else {
            u.load(p, s_fb, e, i < c.Fq);
        }
        const int cf = __popc(__ballot_sync(0xffffffffu, u.fl & 1)), ce = __popc(__ballot_sync(0xffffffffu, u.fl & 2));
        scanned += (unsigned long long)__popc(__ballot_sync(0xffffffffu, u.has));
        if (lane == 0) {
            const unsigned long long slot = c.slot0 + tile;
            p.tcount[slot] = make_uint2((unsigned)cf, (unsigned)ce);
            if (cf) atomicAdd(&p.gsum[slot / kGroup].x, (unsigned long long)cf);
            if (ce) atomicAdd(&p.gsum[slot / kGroup].y, (unsigned long long)ce);
        }
    }
    if (lane == 0 && scanned) atomicAdd(&p.ctl->E[p.s], scanned);
}

template <int Q, bool WIDE = false>
__global__ void __launch_bounds__(kThreads, RCA_MINB_UNIT) fill_unit(const __grid_constant__ LevelParams p) {
    extern __shared__ int s_dyn[];                    // per warp: 32 finished rows
    __shared__ unsigned long long s_fb[kLv];
    if (launch_void(p)) return;
    UnitCtx c = unit_setup(p, s_fb);
    c.fb_s += p.ctl->subF;
    c.ob_s += p.ctl->subN * (unsigned long long)p.row_len;
    constexpr int S = Q + 1;                          // hops of the rows
    const int lane = threadIdx.x & 31, rl = p.row_len;
    const unsigned lt = (1u << lane) - 1u;
    int *stage = s_dyn + (threadIdx.x >> 5) * (kTile * rl);
    const bool want_rows = !p.count_only && (p.unit_fl & 2);
    const unsigned nwarps = gridDim.x * kWarps;
    const int4 *fq = p.frontier + s_fb[Q];            // wide: plane A
    const unsigned long long wstride = WIDE ? wide_stride(c.Fq) : 0ull;
    const int4 none = make_int4(0, 0, 0, 0);
    unsigned tile = blockIdx.x * kWarps + (threadIdx.x >> 5);
    int4 e_next = ((unsigned long long)tile * kTile + lane < c.Fq) ? fq[(unsigned long long)tile * kTile + lane] : none;
    int4 b_next = (WIDE && (unsigned long long)tile * kTile + lane < c.Fq) ? fq[wstride + (unsigned long long)tile * kTile + lane] : none;
    for (; tile < c.ntiles; tile += nwarps) {
        const unsigned long long slot = c.slot0 + tile, g = slot / kGroup;
        const unsigned in_g = (unsigned)(slot % kGroup);
        uint2 mine = make_uint2(0, 0);
        if ((unsigned)lane < in_g) mine = p.tcount[g * kGroup + lane];
        const ulonglong2 gp = p.gscan[g];
        const unsigned long long i = (unsigned long long)tile * kTile + lane, i2 = i + (unsigned long long)nwarps * kTile;
        const bool valid = i < c.Fq;
        const int4 e = e_next;
        e_next = i2 < c.Fq ? fq[i2] : none;           // the next tile's entries travel while this tile is worked on
        UnitLane<Q> u;
        if constexpr (WIDE) {
            const int4 B = b_next;
            b_next = i2 < c.Fq ? fq[wstride + i2] : none;
            u.load_wide(p, e, B, valid);
        } else {
            u.load(p, s_fb, e, valid);
        }
        // state probes: the header sector of every node of the row, all in flight together (a wide entry names all the
        // nodes at once: the headers leave in the same round trip as the next-hop load)
        const bool probe = want_rows && p.with_states && u.has && (!p.need_inter || u.inter);
        long long ts = 0;
        const States st{p.st_head, p.n_head, p.st_off, p.st_rec, p.st_more};
        ulonglong2 h[S + 1];
        if (p.with_states) {
            if (probe) ts = __ldg(p.seed_ts + u.root);
#pragma unroll
            for (int d = 0; d <= S; d++) h[d] = ld_head(st, (d <= Q) ? u.a[d] : u.v, probe);
        }
        const unsigned long long run_f = gp.x + __reduce_add_sync(0xffffffffu, mine.x);
        const unsigned long long run_e = gp.y + __reduce_add_sync(0xffffffffu, mine.y);
        const unsigned bf = __ballot_sync(0xffffffffu, u.fl & 1), be = __ballot_sync(0xffffffffu, u.fl & 2);
        if (u.fl & 1) {
            const unsigned long long pos = c.fb_s + run_f + __popc(bf & lt);
            if (pos < p.fcap) p.frontier[pos] = make_int4(u.v, (int)i, u.r, p.unit_aux | (u.inter << 16));
        }
        if (!be || p.count_only) continue;
        const int n = __popc(be);
        const unsigned long long w0 = c.ob_s + run_e * (unsigned long long)rl;
        if (w0 + (unsigned long long)n * rl > p.ocap) continue;      // flagged with the counts: the run is void
        if constexpr (S == 3) {
            // Three-hop rows with state columns are 12 words (8 with the mask alone): whole 16-byte words.  When the
            // level's rows start on a 16-byte boundary every lane stores its row straight from registers with two or
            // three vector stores - no staging through shared memory (12 stores with 4-way bank conflicts, 12 loads, 12
            // coalesced stores and a loop per tile; the L1 data pipe, which all of those go through, is what bounds
            // this kernel).
            if (p.with_states && ((w0 & 3ull) == 0)) {
                if (u.fl & 2) {
                    int sid[4];
                    unsigned miss = 0;
#pragma unroll
                    for (int d = 0; d <= S; d++) {
                        int cnt;
                        head_finish(st, h[d], (d <= Q) ? u.a[d] : u.v, ts, sid[d], cnt);
                        miss |= (cnt == 0 ? 1u : 0u) << d;
                    }
                    int4 *dst4 = reinterpret_cast<int4 *>(p.out + w0 + (unsigned long long)__popc(be & lt) * rl);
                    dst4[0] = make_int4(u.a[0], u.a[1], u.a[2], u.v);
                    if (p.with_states == 1) {
                        dst4[1] = make_int4(u.rel[1], u.rel[2], u.r, sid[0]);
                        dst4[2] = make_int4(sid[1], sid[2], sid[3], (int)miss);
                    } else {
                        dst4[1] = make_int4(u.rel[1], u.rel[2], u.r, (int)miss);
                    }
                }
                continue;
            }
        }
        if (u.fl & 2) {
            int *row = stage + __popc(be & lt) * rl;
#pragma unroll
            for (int d = 0; d <= Q; d++) row[d] = u.a[d];
            row[S] = u.v;
#pragma unroll
            for (int d = 1; d <= Q; d++) row[S + d] = u.rel[d];
            row[2 * S] = u.r;
            if (p.with_states) {
                unsigned miss = 0;
#pragma unroll
                for (int d = 0; d <= S; d++) {
                    int first, cnt;
                    head_finish(st, h[d], (d <= Q) ? u.a[d] : u.v, ts, first, cnt);
                    if (p.with_states == 1) row[2 * S + 1 + d] = first;
                    miss |= (cnt == 0 ? 1u : 0u) << d;
                }
                row[p.with_states == 1 ? 3 * S + 2 : 2 * S + 1] = (int)miss;
            }
        }
        __syncwarp();
        int *dst = p.out + w0;
        for (int k = lane; k < n * rl; k += 32) dst[k] = stage[k];
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------
// fused path: count, decoupled look-back, fill in one launch (small seed sets)

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
    const unsigned ep = epoch & 0x3fffffffu;
    while (true) {
        const int t = base - (int)lane;
        unsigned status = kStPre;                           // virtual zero prefix before tile 0
        unsigned long long f = 0, e = 0;
        if (t >= 0) {
            while (true) {
                const ulonglong2 w = ld_state(ts + t);
                status = (unsigned)(w.x >> 62);
                if (status != 0 && ((unsigned)(w.x >> 32) & 0x3fffffffu) == ep) { f = w.x & 0xffffffffull; e = w.y; break; }
                __nanosleep(64);                            // not published yet
            }
        }
        // lanes nearer than the first published prefix hold tile aggregates (each far below 2^32,
        // so 32-bit warp adds are exact); that prefix itself is 64-bit and comes by broadcast
        const unsigned m = __ballot_sync(0xffffffffu, status == kStPre);
        const int first = m ? (__ffs(m) - 1) : 32;
        const bool is_agg = (int)lane < first;
        run_f += __reduce_add_sync(0xffffffffu, is_agg ? (unsigned)f : 0u);
        run_e += __reduce_add_sync(0xffffffffu, is_agg ? (unsigned)e : 0u);
        if (m) {
            run_f += __shfl_sync(0xffffffffu, f, first);
            run_e += __shfl_sync(0xffffffffu, e, first);
            break;
        }
        base -= 32;
    }
    pre_f = run_f;
    pre_e = run_e;
    if (lane == 0) st_state(ts + tile, st_pack(kStPre, epoch, run_f + agg_f), run_e + agg_e);
}

template <bool ADJ2, int MODE, int QT>
__global__ void __launch_bounds__(kThreads, RCA_MINB) fused_level(const __grid_constant__ LevelParams p) {
    extern __shared__ int s_dyn[];
    __shared__ TablesSm tsm;
    __shared__ unsigned long long s_fb[kLv];
    Ctl *ctl = p.ctl;
    // an arena overflowed at an earlier level: entries past its end were never written, the host
    // will grow the arena and run the plan again (checked before any ticket is taken)
    if (launch_void(p)) return;
    const LevelCtx c = level_setup(p, tsm, s_fb, s_dyn);
    const int s = p.s, lane = c.w.lane;
    ulonglong2 *tstate = p.tstate + c.slot0;
    const unsigned epoch = p.epoch ? p.epoch : ctl->epoch;
    unsigned long long scanned = 0;
    unsigned ov = 0;
    while (true) {
        // one ticket per tile, taken only when the warp is ready for it: a tile claimed ahead of
        // time would publish its counts late and stall the look-back of every later tile
        // (measured: prefetching the next ticket made the kernel 4x slower)
        unsigned tile = 0;
        if (lane == 0) tile = atomicAdd(&ctl->ticket[s], 1u);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= c.ntiles) break;
        const int C = tile_prologue<MODE>(p, c.T, c.w, tile, c.Fq, s_fb);
        scanned += (unsigned long long)C;
        int cf, ce, c2, s2;
        tile_count<ADJ2, MODE, QT>(p, c.T, c.w, C, cf, ce, c2, s2);
        unsigned long long run_f, run_e;
        tile_lookback(tstate, tile, epoch, (unsigned long long)cf, (unsigned long long)ce, run_f, run_e);
        if (run_f + cf > 0xffffffffull) ov |= 1u;     // the 32-bit count field of the state word
        if (c.fb_s + run_f + cf > p.fcap) ov |= 1u;
        if (!p.count_only && c.ob_s + (run_e + ce) * (unsigned long long)p.row_len > p.ocap) ov |= 2u;
        if (tile == c.ntiles - 1 && lane == 0) {      // the last tile knows the level totals
            ctl->F[s] = run_f + cf;
            ctl->N[s] = run_e + ce;
            if (p.shortest_only && run_e + ce > 0) ctl->stop_level = (unsigned)s;
        }
        __syncwarp();
        tile_fill<ADJ2, MODE, QT>(p, c.T, c.w, tile, C, c.fb_s, c.ob_s, run_f, run_e);
    }
    if (lane == 0) {
        if (scanned) atomicAdd(&ctl->E[s], scanned);
        if (ov) atomicOr(&ctl->overflow, ov);
    }
}

// ------------------------------------------------------------------------------------------------
// chain_walk: a whole typed chain (generate_query/generate_query.py:241-260) in ONE launch, for chains
// whose steps fan out little (the engine checks the per-edge-type maximum degrees).  One thread per
// seed walks the steps depth-first twice: once to count its rows, once - after a warp scan and the
// same decoupled look-back the fused kernel uses - to write them.  Rows come out ordered by seed and
// then by adjacency slot at each step, which is exactly the order of the level-wise kernels.
struct ChainParams {
    int L;
    const int2 *rng[RCA_MAX_HOPS];    // typed (begin, end) table of step i, indexed by node - row_lo[i]
    const int2 *unit[RCA_MAX_HOPS];   // non-null: no node has two relationships of step i's type on the near side;
                                      // the table holds the (neighbour, relationship) itself, or (-1, -1)
    int row_lo[RCA_MAX_HOPS], row_hi[RCA_MAX_HOPS];
    int against[RCA_MAX_HOPS];        // step walks in_adj (neighbour, rel) instead of out_dst
    const int *adj1;
    const int2 *adj2;
    int n_nodes;
    int with_states, count_only, row_len;
    int one_ticket;                   // every CTA takes ONE tile and leaves (the grid has a CTA per tile): SM slots come
                                      // free all the time, so a kernel of another stream can move in beside this one
    unsigned epoch;
    const int *seeds;
    const long long *seed_ts;
    long long n_seeds;
    int *out;
    unsigned long long ocap;
    Ctl *ctl;
    ulonglong2 *tstate;
    const StHead *st_head;
    const StMore *st_more;
    int n_head;
    const int *st_off;
    const StRec *st_rec;
};

template <bool FILL>
__device__ __forceinline__ unsigned long long chain_dfs(const ChainParams &p, int seed, long long ts, int *dst,
                                                        unsigned long long *scanned, unsigned long long *alive) {
    int node[RCA_MAX_HOPS + 1], rel[RCA_MAX_HOPS], cur[RCA_MAX_HOPS], end[RCA_MAX_HOPS];
    int uv[RCA_MAX_HOPS], ur[RCA_MAX_HOPS];           // unit steps: the one candidate itself
    unsigned long long rows = 0;
    const int L = p.L;
    node[0] = seed;
    // candidates of step d for a path whose tail is v
    auto enter = [&](int d, int v) {
        cur[d] = end[d] = 0;
        if (v < p.row_lo[d] || v >= p.row_hi[d]) return;
        if (p.unit[d]) {
            const int2 nx = __ldg(p.unit[d] + (v - p.row_lo[d]));
            uv[d] = nx.x; ur[d] = nx.y;
            end[d] = nx.x >= 0 ? 1 : 0;
        } else {
            const int2 rg = __ldg(p.rng[d] + (v - p.row_lo[d]));
            cur[d] = rg.x; end[d] = rg.y;
        }
        if (!FILL) scanned[d + 1] += (unsigned long long)(end[d] - cur[d]);
    };
    enter(0, seed);
    int d = 0;
    while (d >= 0) {
        if (cur[d] >= end[d]) { d--; continue; }
        const int slot = cur[d]++;
        int v, r;
        if (p.unit[d]) {
            v = uv[d]; r = ur[d];
        } else if (p.against[d]) {
            const int2 a = __ldg(p.adj2 + slot);
            v = a.x; r = a.y & kRelMask;
        } else {
            v = __ldg(p.adj1 + slot);
            r = slot;
        }
        node[d + 1] = v; rel[d] = r;
        if (!FILL) alive[d + 1]++;
        if (d + 1 == L) {
            if (FILL) {
                int *row = dst + rows * (unsigned long long)p.row_len;
                for (int k = 0; k <= L; k++) row[k] = node[k];
                for (int k = 0; k < L; k++) row[L + 1 + k] = rel[k];
                if (p.with_states) {
                    // every node's header load goes out before the first one is looked at
                    const States st{p.st_head, p.n_head, p.st_off, p.st_rec, p.st_more};
                    ulonglong2 h[RCA_MAX_HOPS + 1];
#pragma unroll
                    for (int k = 0; k <= RCA_MAX_HOPS; k++) h[k] = ld_head(st, k <= L ? node[k] : 0, k <= L);
                    unsigned miss = 0;
#pragma unroll
                    for (int k = 0; k <= RCA_MAX_HOPS; k++) {
                        if (k > L) continue;
                        int first, cnt;
                        head_finish(st, h[k], node[k], ts, first, cnt);
                        if (p.with_states == 1) row[2 * L + 1 + k] = first;
                        miss |= (cnt == 0 ? 1u : 0u) << k;
                    }
                    row[p.with_states == 1 ? 3 * L + 2 : 2 * L + 1] = (int)miss;
                }
            }
            rows++;
            continue;
        }
        d++;
        enter(d, v);
    }
    return rows;
}

__device__ __forceinline__ void chain_walk_body(const ChainParams &p) {
    // One ticket = 256 seeds (one per thread).  Row counts are scanned inside the CTA first and only the
    // CTA totals go through the look-back: every tile of a call is in flight at once (100 k seeds fill a
    // third of the GPU), so a tile waits for (tiles before it) / 32 look-back steps - 8 x fewer this way.
    __shared__ unsigned s_tile;
    __shared__ unsigned long long s_wsum[kWarps], s_pre;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    Ctl *ctl = p.ctl;
    const unsigned epoch = p.epoch ? p.epoch : ctl->epoch;
    const unsigned ntiles = (unsigned)((p.n_seeds + kThreads - 1) / kThreads);
    const int L = p.L;
    unsigned long long sc[RCA_MAX_HOPS + 1], al[RCA_MAX_HOPS + 1];
    for (int k = 0; k <= RCA_MAX_HOPS; k++) { sc[k] = 0; al[k] = 0; }
    bool ov = false;
    for (bool once = true; once || !p.one_ticket; once = false) {
        __syncthreads();                              // the previous tile's readers of s_* are done
        if (threadIdx.x == 0) s_tile = atomicAdd(&ctl->ticket[L], 1u);
        __syncthreads();
        const unsigned tile = s_tile;
        if (tile >= ntiles) break;
        const long long i = (long long)tile * kThreads + threadIdx.x;
        int seed = -1;
        long long ts = 0;
        if (i < p.n_seeds) {
            seed = __ldg(p.seeds + i);
            if (p.with_states) ts = __ldg(p.seed_ts + i);
        }
        unsigned long long mine = 0;
        if (seed >= 0 && seed < p.n_nodes) mine = chain_dfs<false>(p, seed, ts, nullptr, sc, al);
        unsigned long long incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        if (lane == 31) s_wsum[wid] = incl;
        __syncthreads();
        unsigned long long tile_rows = 0, before = 0;     // rows of the whole tile / of the warps before this one
#pragma unroll
        for (int k = 0; k < kWarps; k++) {
            const unsigned long long x = s_wsum[k];
            tile_rows += x;
            if (k < wid) before += x;
        }
        if (wid == 0) {
            unsigned long long pre_f, pre_e;
            tile_lookback(p.tstate, tile, epoch, 0ull, tile_rows, pre_f, pre_e);
            if (lane == 0) {
                s_pre = pre_e;
                if (tile == ntiles - 1) ctl->N[L] = pre_e + tile_rows;
            }
        }
        __syncthreads();
        const unsigned long long pre = s_pre;
        const unsigned long long base = ctl->out_base + (pre + before + incl - mine) * (unsigned long long)p.row_len;
        if (!p.count_only && ctl->out_base + (pre + tile_rows) * (unsigned long long)p.row_len > p.ocap) ov = true;
        else if (!p.count_only && mine) chain_dfs<true>(p, seed, ts, p.out + base, nullptr, nullptr);
    }
    // counters: partial paths alive after each step and entries scanned at each step
    for (int k = 1; k <= L; k++) {
        unsigned long long a = al[k], b = sc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
        if (lane == 0) {
            if (a && k < L) atomicAdd(&ctl->F[k], a);
            if (b) atomicAdd(&ctl->E[k], b);
        }
    }
    if (ov && lane == 0) atomicOr(&ctl->overflow, 2u);
}

// chain_line: the same job as chain_walk for chains whose steps before the last are all unit steps (the
// usual case: EVENT <-HasEvent- Event -involvedObject-> Pod -spec_nodeName-> Node).  A seed then has ONE
// partial path, so there is nothing to search: the walk is a straight line of single loads, nodes and
// relationships stay in registers across the look-back (no second walk, no local-memory stacks), the
// state headers of the nodes are requested as the walk passes them, and only the last step may fan out
// into several rows.  Same rows, same order, same counters as chain_walk.
__device__ __forceinline__ void chain_line_body(const ChainParams &p) {
    __shared__ unsigned s_tile;
    __shared__ unsigned long long s_wsum[kWarps], s_pre;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    Ctl *ctl = p.ctl;
    const unsigned epoch = p.epoch ? p.epoch : ctl->epoch;
    const unsigned ntiles = (unsigned)((p.n_seeds + kThreads - 1) / kThreads);
    const int L = p.L;
    const bool probe = p.with_states && !p.count_only;
    const States st{p.st_head, p.n_head, p.st_off, p.st_rec, p.st_more};
    unsigned sc[RCA_MAX_HOPS + 1], al[RCA_MAX_HOPS + 1];
#pragma unroll
    for (int k = 0; k <= RCA_MAX_HOPS; k++) { sc[k] = 0; al[k] = 0; }
    bool ov = false;
    for (bool once = true; once || !p.one_ticket; once = false) {
        __syncthreads();                              // the previous tile's readers of s_* are done
        if (threadIdx.x == 0) s_tile = atomicAdd(&ctl->ticket[L], 1u);
        __syncthreads();
        const unsigned tile = s_tile;
        if (tile >= ntiles) break;
        const long long i = (long long)tile * kThreads + threadIdx.x;
        int seed = -1;
        long long ts = 0;
        if (i < p.n_seeds) {
            seed = __ldg(p.seeds + i);
            if (p.with_states) ts = __ldg(p.seed_ts + i);
        }
        int node[RCA_MAX_HOPS + 1], rel[RCA_MAX_HOPS];
        ulonglong2 h[RCA_MAX_HOPS];
        bool alive = seed >= 0 && seed < p.n_nodes;
        node[0] = seed;
        int lb = 0, le = 0;                           // candidates of the last step
        int2 last = make_int2(-1, -1);                // ... or the one candidate itself, if that step is a unit step too
#pragma unroll
        for (int k = 0; k < RCA_MAX_HOPS; k++) {
            h[k] = make_ulonglong2(1ull, 0ull);
            if (k >= L) continue;
            if (probe) h[k] = ld_head(st, node[k], alive);
            const bool in = alive && node[k] >= p.row_lo[k] && node[k] < p.row_hi[k];
            if (k < L - 1 || p.unit[k]) {
                int2 nx = make_int2(-1, -1);
                if (in) nx = __ldg(p.unit[k] + (node[k] - p.row_lo[k]));
                const bool ok = nx.x >= 0;
                sc[k + 1] += ok ? 1u : 0u;
                if (k < L - 1) {
                    alive = ok;
                    node[k + 1] = nx.x; rel[k] = nx.y;
                    al[k + 1] += ok ? 1u : 0u;
                } else {
                    last = nx;
                    le = ok ? 1 : 0;
                }
            } else if (in) {
                const int2 rg = __ldg(p.rng[k] + (node[k] - p.row_lo[k]));
                lb = rg.x; le = rg.y;
                sc[k + 1] += (unsigned)(rg.y - rg.x);
            }
        }
        const unsigned long long mine = (unsigned long long)(le - lb);
        unsigned long long incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        if (lane == 31) s_wsum[wid] = incl;
        __syncthreads();
        unsigned long long tile_rows = 0, before = 0;
#pragma unroll
        for (int k = 0; k < kWarps; k++) {
            const unsigned long long x = s_wsum[k];
            tile_rows += x;
            if (k < wid) before += x;
        }
        if (wid == 0) {
            unsigned long long pre_f, pre_e;
            tile_lookback(p.tstate, tile, epoch, 0ull, tile_rows, pre_f, pre_e);
            if (lane == 0) {
                s_pre = pre_e;
                if (tile == ntiles - 1) ctl->N[L] = pre_e + tile_rows;
            }
        }
        __syncthreads();
        const unsigned long long pre = s_pre;
        if (p.count_only) continue;
        if (ctl->out_base + (pre + tile_rows) * (unsigned long long)p.row_len > p.ocap) { ov = true; continue; }
        if (!mine) continue;
        // the states of the nodes every row of this seed shares
        int first[RCA_MAX_HOPS];
        unsigned miss0 = 0;
        if (probe) {
#pragma unroll
            for (int k = 0; k < RCA_MAX_HOPS; k++) {
                first[k] = -1;
                if (k >= L) continue;
                int cnt;
                head_finish(st, h[k], node[k], ts, first[k], cnt);
                miss0 |= (cnt == 0 ? 1u : 0u) << k;
            }
        }
        int *row = p.out + ctl->out_base + (pre + before + incl - mine) * (unsigned long long)p.row_len;
        for (int j = lb; j < le; j++, row += p.row_len) {
            int v, r;
            if (p.unit[L - 1]) {
                v = last.x; r = last.y;
            } else if (p.against[L - 1]) {
                const int2 a = __ldg(p.adj2 + j);
                v = a.x; r = a.y & kRelMask;
            } else {
                v = __ldg(p.adj1 + j);
                r = j;
            }
#pragma unroll
            for (int k = 0; k < RCA_MAX_HOPS; k++) {
                if (k >= L) continue;
                row[k] = node[k];
                if (k < L - 1) row[L + 1 + k] = rel[k];
                if (probe && p.with_states == 1) row[2 * L + 1 + k] = first[k];
            }
            row[L] = v;
            row[2 * L] = r;
            if (probe) {
                int fv, cnt;
                state_probe(st, v, ts, fv, cnt);
                if (p.with_states == 1) row[3 * L + 1] = fv;
                row[p.with_states == 1 ? 3 * L + 2 : 2 * L + 1] = (int)(miss0 | ((cnt == 0 ? 1u : 0u) << L));
            }
        }
    }
    for (int k = 1; k <= L; k++) {
        unsigned a = al[k], b = sc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
        if (lane == 0) {
            if (a && k < L) atomicAdd(&ctl->F[k], (unsigned long long)a);
            if (b) atomicAdd(&ctl->E[k], (unsigned long long)b);
        }
    }
    if (ov && lane == 0) atomicOr(&ctl->overflow, 2u);
}

__global__ void __launch_bounds__(kThreads) chain_walk(const __grid_constant__ ChainParams p) { chain_walk_body(p); }
__global__ void __launch_bounds__(kThreads) chain_line(const __grid_constant__ ChainParams p) { chain_line_body(p); }

// Several one-launch chains in ONE launch (rca_expand_many: the instance queries of a submission): blockIdx.y
// picks the chain.  Five chains of 100 k seeds each fill the GPU together where each alone fills a third of it,
// and the host enqueues one kernel instead of five.  CTAs beyond a chain's own tile count leave at their first
// ticket.
constexpr int kChainBatch = 8;
struct ChainBatch {
    ChainParams p[kChainBatch];
    int line[kChainBatch];            // 1: chain_line, 0: chain_walk
};
static_assert(sizeof(ChainBatch) <= 4000, "ChainBatch must fit the kernel parameter space");
#ifndef RCA_MINB_BATCH
#define RCA_MINB_BATCH 2
#endif
__global__ void __launch_bounds__(kThreads, RCA_MINB_BATCH) chain_batch(const __grid_constant__ ChainBatch b) {
    const ChainParams &p = b.p[blockIdx.y];
    if (b.line[blockIdx.y]) chain_line_body(p);
    else chain_walk_body(p);
}

// rca_expand_many: the next plan's rows start where this plan's end (row_len 0 for count-only plans)
struct RowLens { int v[RCA_MAX_HOPS + 1]; };
__global__ void chain_plans(const Ctl *prev, Ctl *next, RowLens rl) {
    if (threadIdx.x == 0) {
        unsigned long long end = prev->out_base;
        for (int j = 1; j <= RCA_MAX_HOPS; j++) end += prev->N[j] * (unsigned long long)rl.v[j];
        next->out_base = end;
        if (prev->overflow) next->overflow = prev->overflow;      // a void run voids what follows it
    }
}

// Stand-alone HasState predicate, one thread per query: intervals of entity nodes[i] with
// tmin <= ta[i] and tmax > tb[i].  ta == tb is the strict form (analyze_root_cause.py:70-79);
// (ta, tb) = (window end, window start) is the loose form (:51-60).
// [st_lo, st_hi): only STATE nodes in this id range count (ids are kind-major, so this is the `(n2:KIND)` label of
// the statement; the reference applies the label before its LIMIT 10).
__global__ void state_lookup(const int *st_off, const StRec *rec, int n_nodes, const int *nodes, const long long *ta,
                             const long long *tb, long long n, int st_lo, int st_hi, int *out_first, int *out_cnt,
                             int *out_ids) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int v = nodes[i], c = 0, f = -1;
    if (v >= 0 && v < n_nodes) {
        const long long a = ta[i], b = tb[i];
        for (int j = st_off[v]; j < st_off[v + 1]; j++) {
            if (rec[j].tmin <= a && rec[j].tmax > b) {
                const int sn = rec[j].node;
                if (sn < st_lo || sn >= st_hi) continue;
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

// ------------------------------------------------------------------------------------------------
// Batched `message CONTAINS pattern` (find_metapath/find_srckind_metapath_neo4j.py:78,
// generate_query/generate_query.py:225): which EVENT messages contain which of P patterns.
// Rabin-Karp: one thread per (message, distinct pattern length) rolls a 64-bit polynomial hash over
// the message's windows of that length and probes an open-addressing table of pattern hashes; a hit
// is verified byte by byte, then recorded: atomicMin of the message index per pattern (the
// canonical 'LIMIT 1' answer) and a (pattern, message) pair appended to a list.
constexpr unsigned long long kRkBase = 0x100000001b3ull;          // odd multiplier (FNV prime)

struct TextParams {
    const unsigned char *msg;         // message bytes, concatenated
    const long long *msg_off;         // [M+1]
    long long n_msgs;
    const unsigned char *pat;         // pattern bytes, concatenated (distinct patterns)
    const long long *pat_off;         // [P+1]
    const int *lens;                  // [D] distinct pattern lengths (> 0)
    const unsigned long long *pow_top;// [D] base^(len-1)
    int n_lens;
    const unsigned long long *tab_hash;   // [T] hash of the pattern in the slot
    const int *tab_pat;               // [T] pattern index or -1 (empty); T is a power of two
    unsigned tab_mask;
    int *first;                       // [P] lowest matching message index (init INT_MAX)
    int2 *pairs;                      // [cap] (pattern, message)
    unsigned long long cap;
    unsigned long long *n_pairs;      // matches found (may exceed cap: then the list is truncated)
};

__device__ __forceinline__ unsigned rk_slot(unsigned long long h, int len, unsigned mask) {
    unsigned long long x = h ^ ((unsigned long long)len * 0x9e3779b97f4a7c15ull);
    x ^= x >> 29; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 32;
    return (unsigned)x & mask;
}

__global__ void __launch_bounds__(256) contains_scan(const __grid_constant__ TextParams p) {
    const long long gid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long m = gid / p.n_lens;
    if (m >= p.n_msgs) return;
    const int d = (int)(gid - m * p.n_lens);
    const int len = p.lens[d];
    const long long b = p.msg_off[m];
    const int n = (int)(p.msg_off[m + 1] - b);
    if (n < len) return;
    const unsigned char *t = p.msg + b;
    const unsigned long long top = p.pow_top[d];
    unsigned long long h = 0;
    for (int i = 0; i < len; i++) h = h * kRkBase + t[i];
    for (int pos = 0;; pos++) {
        for (unsigned sl = rk_slot(h, len, p.tab_mask);; sl = (sl + 1) & p.tab_mask) {
            const int pi = p.tab_pat[sl];
            if (pi < 0) break;
            if (p.tab_hash[sl] != h) continue;
            const long long pb = p.pat_off[pi];
            if ((int)(p.pat_off[pi + 1] - pb) != len) continue;
            bool eq = true;
            for (int i = 0; i < len && eq; i++) eq = p.pat[pb + i] == t[pos + i];
            if (eq) {
                atomicMin(p.first + pi, (int)m);
                const unsigned long long k = atomicAdd(p.n_pairs, 1ull);
                if (k < p.cap) p.pairs[k] = make_int2(pi, (int)m);
            }
        }
        if (pos + len >= n) break;
        h = (h - t[pos] * top) * kRkBase + t[pos + len];
    }
}


// ------------------------------------------------------------------------------------------------
// match_pairs: the message filter that follows an instance query (generate_query/generate_query.py:104-129:
// a row survives iff its destination's name, or its kind, occurs in the event's message), for many
// (message, name, kind) triples at once.  One warp per triple: the lanes take the start positions of the
// message 32 at a time and compare bytes; messages are a hundred bytes, needles ten or twenty.
struct PairParams {
    const unsigned char *msg;         // the loaded message arena (rca_text_load)
    const long long *msg_off;
    const unsigned char *ndl;         // needle bytes, concatenated
    const long long *ndl_off;
    const int *hay, *na, *nb;         // [n] message index, first and second needle index (-1: none)
    long long n;
    unsigned char *out;               // [n] 1 = one of the needles occurs in the message
};

__device__ __forceinline__ bool warp_find(const unsigned char *h, int hn, const unsigned char *x, int xn, int lane) {
    if (xn == 0) return true;                         // '' in s
    if (xn > hn) return false;
    const unsigned char x0 = x[0];
    for (int base = 0; base <= hn - xn; base += 32) {
        const int pos = base + lane;
        bool ok = pos <= hn - xn && h[pos] == x0;
        for (int k = 1; ok && k < xn; k++) ok = h[pos + k] == x[k];
        if (__any_sync(0xffffffffu, ok)) return true;
    }
    return false;
}

__global__ void __launch_bounds__(256) match_pairs(const __grid_constant__ PairParams p) {
    const long long k = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (k >= p.n) return;
    const int m = p.hay[k];
    const long long hb = p.msg_off[m];
    const int hn = (int)(p.msg_off[m + 1] - hb);
    bool hit = false;
    const int a = p.na[k], b = p.nb[k];
    if (a >= 0) hit = warp_find(p.msg + hb, hn, p.ndl + p.ndl_off[a], (int)(p.ndl_off[a + 1] - p.ndl_off[a]), lane);
    if (!hit && b >= 0) hit = warp_find(p.msg + hb, hn, p.ndl + p.ndl_off[b], (int)(p.ndl_off[b + 1] - p.ndl_off[b]), lane);
    if (lane == 0) p.out[k] = hit ? 1 : 0;
}


// ------------------------------------------------------------------------------------------------
// peer_exchange: the one collective of the path - every rank learns every rank's instance count
//  - done by the ranks themselves over NVLink / NVSwitch peer memory.  Every engine
// owns an array of `world` slots {value, seq}; rank r's thread t stores its value, then the sequence
// number, into slot r of rank t's array (a plain remote store and a release store), and then waits
// until slot t of its OWN array carries the same sequence number (or, deferred, the previous one).  One launch
// of `world` threads, no host involvement, no NCCL call on the step's critical path; the wait gives up after
// about two seconds and raises *err instead of hanging the GPU.
struct PeerSlot {
    long long value;
    unsigned long long seq;
};
struct PeerParams {
    PeerSlot *slots[8];               // slots[t]: rank t's array of this exchange's buffer (own array at index rank)
    PeerSlot *prev[8];                // ... and of the previous exchange's buffer (deferred mode)
    int rank, world;
    unsigned long long seq;           // this exchange's number (1, 2, ...; the same on every rank)
    int publish;                      // 0: only wait for / collect exchange `seq` (rca_peer_wait)
    int deferred;                     // 1: publish `seq`, but wait for and collect exchange seq-1: the ranks are then
                                      //    coupled one step apart instead of step by step (no per-step wait for the
                                      //    slowest rank); the last exchange is collected by rca_peer_wait
    long long host_value;             // value to publish
    long long *out;                   // [world] the collected values, rank order (device memory of this rank)
    unsigned *err;
    unsigned long long timeout_ns;    // give up waiting for a peer after this long (rca_peer_set_timeout)
};
// Four buffers rotate (seq & 3).  A rank can run ahead of another by at most one finished exchange plus the one in
// flight; with the deferred wait that is three buffers in use at once, so the fourth write never lands on values a
// slow rank has yet to read.
constexpr int kPeerBuffers = 4;

__global__ void peer_exchange(const __grid_constant__ PeerParams p) {
    const int t = threadIdx.x;
    if (t >= p.world) return;
    if (p.publish) {
        PeerSlot *dst = p.slots[t] + p.rank;
        dst->value = p.host_value;
        asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(&dst->seq), "l"(p.seq) : "memory");
    }
    const unsigned long long want = p.deferred ? p.seq - 1 : p.seq;
    if (want == 0) { p.out[t] = 0; return; }
    const PeerSlot *src = (p.deferred ? p.prev[p.rank] : p.slots[p.rank]) + t;
    unsigned long long t0 = 0, now = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (true) {
        unsigned long long s;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(s) : "l"(&src->seq) : "memory");
        if (s >= want) break;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (now - t0 > p.timeout_ns) { atomicOr(p.err, 1u); p.out[t] = -1; return; }     // the slot was never written
        __nanosleep(200);
    }
    p.out[t] = src->value;
}

// ------------------------------------------------------------------------------------------------
// push_rows: the payload half of the gather ("gather-v of compacted instances").  Every
// rank's rows of one level are already compact in its own output arena; after the count exchange each rank
// knows where its rows start in the destination rank's window (exclusive scan over the ranks' counts), and
// this kernel stores them there over NVLink peer memory - plain 16-byte remote stores, each thread a
// contiguous run, no staging copy, no host in the data path.  Loads are local and may be unaligned against
// the destination (the two offsets are unrelated), so the stores are aligned on the DESTINATION and each one
// is assembled from four word loads of the local arena.  Completion is published by the peer_exchange launch
// that follows on the same stream (release store after this kernel has finished).
struct PushParams {
    const int *src;                   // first word of this rank's rows (local output arena)
    int *dst;                         // where they go in the destination window (peer or local memory)
    unsigned long long words;
};

__global__ void __launch_bounds__(256) push_rows(const __grid_constant__ PushParams p) {
    const unsigned long long n = p.words;
    // head: words before the first 16-byte boundary of dst
    unsigned long long head = ((16u - (unsigned)(reinterpret_cast<uintptr_t>(p.dst) & 15u)) & 15u) >> 2;
    if (head > n) head = n;
    const unsigned long long vec = (n - head) >> 2, tail0 = head + (vec << 2);
    const unsigned long long tid = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
    const unsigned long long nthreads = gridDim.x * (unsigned long long)blockDim.x;
    if (tid < head) p.dst[tid] = __ldg(p.src + tid);
    if (tid < n - tail0) p.dst[tail0 + tid] = __ldg(p.src + tail0 + tid);
    int4 *d4 = reinterpret_cast<int4 *>(p.dst + head);
    const int *s = p.src + head;
    const bool src_aligned = (reinterpret_cast<uintptr_t>(s) & 15u) == 0;
    for (unsigned long long j = tid; j < vec; j += nthreads) {
        int4 v;
        if (src_aligned) v = __ldg(reinterpret_cast<const int4 *>(s) + j);
        else v = make_int4(__ldg(s + 4 * j), __ldg(s + 4 * j + 1), __ldg(s + 4 * j + 2), __ldg(s + 4 * j + 3));
        d4[j] = v;
    }
    __threadfence_system();
}

}  // namespace rca
