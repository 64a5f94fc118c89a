/* rca_b200.h - C ABI of the B200 path engine (librca_b200.so).
 *
 * This is the boundary a maintainer of freiris/k8s-llm-rca binds instead of the bolt connection
 * to Neo4j.  In the reference every graph operation is
 *     Neo4jQueryExecutor.run_query(query, parameters)      common/neo4j_query_executor.py:15-24
 * which ships Cypher text to a Neo4j server.  Here the Cypher text is translated to a plan on
 * the host (Python, cold path) and the plan is executed by the entry points below against a
 * CSR / typed-edge snapshot resident in HBM.  Plain pointers and sizes only; integer status
 * codes; no C++ exceptions cross this boundary; INTEGRATION.md shows the ctypes stub.
 *
 * Status codes: 0 = ok, <0 = error (rca_last_error gives the text).
 */
#ifndef RCA_B200_H
#define RCA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RCA_MAX_HOPS 8          /* longest path (relationships) a plan may ask for            */
#define RCA_MAX_KINDS 256       /* kinds (entity + state) a snapshot may hold                  */
#define RCA_STATE_LIMIT 10      /* 'LIMIT 10' of the state query, analyze_root_cause.py:77     */

#define RCA_OK 0
#define RCA_ERR_ARG (-1)
#define RCA_ERR_CUDA (-2)
#define RCA_ERR_NOMEM (-3)
#define RCA_ERR_NODEVICE (-4)

/* Direction of one hop relative to the walk. */
#define RCA_DIR_OUT 0           /* (a)-[r]->(b), a bound                                       */
#define RCA_DIR_IN 1            /* (b)<-[r]-(a), a bound: the reference's "reversed" steps,
                                   test_generate_query.py:38                                   */
#define RCA_DIR_UND 2           /* (a)-[r]-(b)                                                 */

#define RCA_STATES_NONE 0
#define RCA_STATES_FULL 1
#define RCA_STATES_MASK 2

#define RCA_PLAN_VARLEN 0       /* find_metapath/find_srckind_metapath_neo4j.py:95-129         */
#define RCA_PLAN_CHAIN 1        /* generate_query/generate_query.py:241-260                    */

typedef struct rca_engine rca_engine_t;

/* Host view of a snapshot (k8s-llm-rca_b200/snapshot.py documents every array).  All pointers
 * are HOST pointers; rca_open copies them to the device and keeps nothing of the host memory. */
typedef struct {
    int32_t n_nodes, n_kinds, n_etypes, reserved;
    int64_t n_rels;                 /* rows of out_dst / in_adj                                  */
    int64_t n_und;                  /* rows of und_adj                                           */
    int64_t n_states;               /* rows of st_*                                              */
    int64_t n_typed_out, n_typed_in;/* rows of typed_out / typed_in                              */
    const int32_t *kind_start;      /* [n_kinds+1]                                               */
    const int32_t *et_src, *et_dst; /* [n_etypes] kind of the start / end node of each edge type */
    const int32_t *et_cls;          /* [n_etypes] 0 entity-entity, 1 touches a STATE kind        */
    const int32_t *out_rng;         /* [n_nodes+1][2] (begin, entity_end)                        */
    const int32_t *out_dst;         /* [n_rels]; slot == relationship id                         */
    const int32_t *in_rng;          /* [n_nodes+1][2]                                            */
    const int32_t *in_adj;          /* [n_rels][2] (src, rel | 1<<30)                            */
    const int32_t *und_rng;         /* [n_nodes+1][2]                                            */
    const int32_t *und_adj;         /* [n_und][2] (neighbour, rel; bit 30 set: hop runs against
                                       the relationship's stored direction)                     */
    const int32_t *typed_out;       /* [n_typed_out][2] (begin, end) into out_dst                */
    const int64_t *typed_out_base;  /* [n_etypes+1]                                              */
    const int32_t *typed_in;        /* [n_typed_in][2] (begin, end) into in_adj                  */
    const int64_t *typed_in_base;   /* [n_etypes+1]                                              */
    const int32_t *st_off;          /* [n_nodes+1] HasState interval lists                       */
    const int64_t *st_tmin, *st_tmax;/* [n_states] milliseconds; interval is [tmin, tmax)        */
    const int32_t *st_node;         /* [n_states] STATE node id                                  */
    const int32_t *rel_et;          /* [n_rels] edge type of every relationship                  */
} rca_snapshot_t;

/* One query plan.  A VARLEN plan is the reference's variable-length metapath query
 * (find_srckind_metapath_neo4j.py:95-116, and :119-129 through pos_kind / node_simple);
 * a CHAIN plan is the per-metapath instance query (generate_query.py:241-260). */
typedef struct {
    int32_t kind;                   /* RCA_PLAN_VARLEN | RCA_PLAN_CHAIN                          */
    /* --- CHAIN --- */
    int32_t n_steps;                /* L, 1..RCA_MAX_HOPS                                        */
    int32_t step_etype[RCA_MAX_HOPS];
    int32_t step_dir[RCA_MAX_HOPS]; /* RCA_DIR_OUT | RCA_DIR_IN                                  */
    /* --- VARLEN --- */
    int32_t direction;              /* RCA_DIR_*                                                 */
    int32_t min_hops, max_hops;     /* '*1..3' -> 1, 3                                           */
    int32_t entity_only;            /* walk only entity-entity relationships                     */
    int32_t node_simple;            /* all(node ... single(...)) : no node repeats (:98)         */
    int32_t dest_kind;              /* n2.kind = $destKind (:97); -1 = any                       */
    int32_t has_inter;              /* size($intermediateKinds) > 0 (:100-102)                   */
    int32_t prune;                  /* drop partial paths that provably cannot reach dest_kind in
                                       the hops that remain: schema reachability, refined by the
                                       edge type a path arrived over (stepping straight back over
                                       an edge type that is functional on that side is impossible).
                                       Result-neutral.                                           */
    int32_t pos_kind[RCA_MAX_HOPS + 1]; /* kind required at path position p, -1 = any (:127)     */
    uint8_t kind_excl[RCA_MAX_KINDS];   /* 1: kind may appear nowhere on a path (:99)            */
    uint8_t kind_inter[RCA_MAX_KINDS];  /* 1: kind is in $intermediateKinds                      */
    /* --- both --- */
    int32_t with_states;            /* fuse the HasState predicate (analyze_root_cause.py:70-79)
                                       into emission.  RCA_STATES_FULL: rows grow by (len+1) state
                                       ids + 1 mask; RCA_STATES_MASK: by the mask alone - bit d set =
                                       node d of the row has no state at the seed's time, which is
                                       what the reference's check branches on (:180) - a third fewer
                                       bytes to carry to the host                                 */
    int32_t count_only;             /* count path instances per length but do not write rows (no
                                       output arena needed; for very large enumerations)         */
    int32_t shortest_only;          /* VARLEN: keep only the rows of the shortest length that has any - the
                                       rule find_metapath applies to what its statement returned
                                       (find_srckind_metapath_neo4j.py:153-154).  The launches of every
                                       longer length return at once: their rows are never made, let
                                       alone carried to the host.  Over several GPUs each rank stops at
                                       ITS shortest length; the caller takes the minimum over the ranks
                                       (parallel.ladder_decision) and drops the longer ranks' rows      */
} rca_plan_t;

/* Counters of the last rca_expand, all exact (they come from the kernels' own count pass). */
typedef struct {
    int64_t frontier[RCA_MAX_HOPS + 1];  /* F_s: partial paths alive after step s (F_0 = seeds)  */
    int64_t scanned[RCA_MAX_HOPS + 1];   /* E_s: adjacency entries scanned at step s             */
    int64_t emitted[RCA_MAX_HOPS + 1];   /* N_s: path instances of length s                      */
    int32_t row_len[RCA_MAX_HOPS + 1];   /* int32 words per row at level s                       */
    int64_t total;                       /* sum of emitted                                       */
    double kernel_ms;                    /* device time of the whole expansion (CUDA events)     */
    double level_ms[RCA_MAX_HOPS + 1];   /* device time of the launch that produced level s      */
    int32_t launches;                    /* kernels launched                                     */
    int32_t retries;                     /* re-runs after growing the scratch arenas             */
} rca_stats_t;

/* replaces GraphDatabase.driver(uri, auth) + verify_connectivity, neo4j_query_executor.py:7-9 */
int rca_open(const rca_snapshot_t *snap, int device, rca_engine_t **out);
/* replaces driver.close(), neo4j_query_executor.py:11-13 */
int rca_close(rca_engine_t *e);
/* text of the last error on this engine (or of the last failed rca_open when e == NULL) */
const char *rca_last_error(const rca_engine_t *e);
/* run the engine's work on this CUDA stream (cudaStream_t as an integer); 0 = its own stream (created
 * with the highest priority).  rca_expand_many runs bounded plans on side streams of the LOWEST priority
 * so that they fill the gaps the main plans leave: a caller's stream should be created above the lowest
 * priority (cudaStreamCreateWithPriority / torch.cuda.Stream(priority=-1)) to get the same effect. */
int rca_set_stream(rca_engine_t *e, uint64_t cuda_stream);
/* Plans with at most `max_seeds` seeds run one fused launch per hop (fewest launches: the
 * per-incident latency case); larger seed sets run count / scan / fill launches per hop (no
 * inter-tile waiting: the throughput case).  Results are identical either way.  Defaults: 16384
 * for variable-length plans, 2^20 for typed chains (whose tiles are uniform, so the fused launch
 * suits them much longer); this call sets both.
 * No counterpart in the reference: it is a tuning knob of this engine. */
int rca_set_fused_threshold(rca_engine_t *e, int64_t max_seeds);
/* Typed chains whose steps fan out little - product of the steps' largest adjacency-row lengths at
 * most `max_rows_per_seed` (default 64; 0 switches this off) - run as ONE depth-first launch instead
 * of one launch per step.  Same rows, same order, same counters.  No counterpart in the reference. */
int rca_set_chain_walk_fanout(rca_engine_t *e, int64_t max_rows_per_seed);
/* A level of a large variable-length plan that comes down to a typed step over an edge type no node
 * has two of on the near side (Pod -spec_nodeName-> Node, ownerReferences, claimRef ...) runs through
 * kernels made for that case: one lane per partial path, no row-length scan, no emission queue, every
 * load of the lane in flight as early as its address is known.  Same rows, same order, same counters.
 * on = 0 sends such levels through the general kernels (default on).  No counterpart in the reference. */
int rca_set_unit_steps(rca_engine_t *e, int on);
/* When the last level of a large variable-length plan consists of such unit steps and the paths before it have at
 * most three nodes, the level before it can store every surviving path whole - nodes, relationships, seed index,
 * 32 bytes in two coalesced planes - instead of a 16-byte parent-pointer entry: the last level then knows all nodes
 * of a row after one contiguous read, where it chases the parent entry and the seed.  Same rows, same order, same
 * counters.  Default OFF: on BASELINE's config 3 the last level gains less (0.189 -> 0.187 ms) than the level
 * before it loses to the wider writes (0.095 -> 0.102 ms); a first version that also carried the state answers lost
 * more (level 2: 0.10 -> 0.18 ms).  No counterpart in the reference. */
int rca_set_wide_entries(rca_engine_t *e, int on);
/* device memory the snapshot occupies on this engine's GPU (adjacency, typed tables, intervals and the
 * tables rca_open derives from them), in bytes.  No counterpart in the reference. */
int64_t rca_graph_bytes(const rca_engine_t *e);
/* pre-size the scratch arenas: frontier entries (16 B each) and output words (4 B each) */
int rca_reserve(rca_engine_t *e, int64_t frontier_entries, int64_t output_words);

/* replaces session.run(query, parameters) for the path-producing statements,
 * neo4j_query_executor.py:16-21.  seeds: node ids (host or device memory, seeds_on_device says
 * which); seed_ts: one timestamp (ms) per seed, needed only with plan->with_states.
 * Results stay in the engine until the next rca_expand / rca_close on it.
 * Device seed buffers (seeds_on_device = 1) are read on the engine's stream, which is not ordered against the
 * stream that wrote them: give the engine that stream (rca_set_stream) or finish the writes first. */
int rca_expand(rca_engine_t *e, const rca_plan_t *plan, const int32_t *seeds, int64_t n_seeds,
               const int64_t *seed_ts, int seeds_on_device, rca_stats_t *stats);
/* Several plans in one submission (no counterpart in the reference, which sends one statement at a
 * time): the plans run back to back on the engine's stream with one upload, one download and ONE
 * synchronisation for the whole set; plan k's rows follow plan k-1's in the output arena.
 * seeds[k] / n_seeds[k] / seed_ts[k] (seed_ts, or its entries for plans without with_states, may be
 * null) are all host or all device memory.  stats (nullable) receives n_plans entries (kernel_ms = the
 * plan's own launches).  rca_result_select chooses the plan
 * that rca_result_level / rca_result_copy address (plan 0 after the call). */
int rca_expand_many(rca_engine_t *e, const rca_plan_t *plans, int32_t n_plans, const int32_t *const *seeds,
                    const int64_t *n_seeds, const int64_t *const *seed_ts, int seeds_on_device, rca_stats_t *stats);
int rca_result_select(rca_engine_t *e, int32_t plan_index);
/* A submission that comes again - the same plans over the same seed buffers (device memory, or pinned host memory)
 * with the engine's arenas unchanged - is captured once as a CUDA graph and from then on replayed with ONE graph
 * launch instead of some twenty-five stream operations (table and control-block uploads, a dozen kernel launches on
 * two streams, timing events, the download): the host side of a 0.5 ms step is 0.2 ms of enqueueing, and under load
 * (one process per GPU, eight to a host) it is what a step waits for.  Same kernels, same results, same statistics.
 * The engine keeps the few most recent graphs; anything that changes what would be enqueued (other plans, seed
 * counts or buffers, a grown arena, a tuning knob) simply runs eagerly and is captured when it repeats.
 * Default OFF: measured on BASELINE's config 3 a replay is 4 % slower than the eager submission on one GPU and
 * 7 % slower with eight processes on one host (0.53 against 0.507 ms, 0.555 against 0.52 ms) - the step is bound
 * by the dependent chain of its kernels, the host keeps ahead of the GPU, and the graph's event and copy nodes cost
 * more than the launches they replace.  It is for hosts that cannot keep ahead.  No counterpart in the reference. */
int rca_set_graphs(rca_engine_t *e, int on);
/* device pointer / shape of the rows of paths with `level` relationships.
 * Row = (level+1) node ids, level relationship ids [, (level+1) state ids][, 1 missing-state mask] */
int rca_result_level(const rca_engine_t *e, int level, const int32_t **dev_rows, int64_t *n_rows,
                     int32_t *row_len);
/* replaces records = list(result), neo4j_query_executor.py:21: copy one level's rows to the host */
int rca_result_copy(rca_engine_t *e, int level, int32_t *host_rows);
/* Overlapping the copy-out with the next submission (no counterpart in the reference, whose driver hands over
 * materialised records).  rca_set_double_buffer(e, 1): submissions alternate between two output arenas.
 * rca_result_copy_async: like rca_result_copy, but only enqueues the copy (host_rows should be pinned) on a
 * copy stream of the engine and returns; the rows keep travelling while the NEXT rca_expand / rca_expand_many
 * runs (it writes the other arena; a submission waits, on the device, for copies that still read the arena it
 * is about to reuse).  rca_result_wait: all enqueued copies have arrived. */
int rca_set_double_buffer(rca_engine_t *e, int on);
int rca_result_copy_async(rca_engine_t *e, int level, int32_t *host_rows);
int rca_result_wait(rca_engine_t *e);

/* replaces the per-node state query of check_states_of_entity, analyze_root_cause.py:173-176:
 * for i < n: STATE nodes of entity nodes[i] whose interval holds ts[i].  out_first[i] = lowest
 * matching STATE node id or -1; out_cnt[i] = number of matches; out_ids (nullable)
 * [n][RCA_STATE_LIMIT], -1 padded.  Buffers are host or device memory per on_device. */
int rca_state_lookup(rca_engine_t *e, const int32_t *nodes, const int64_t *ts, int64_t n, int on_device,
                     int32_t *out_first, int32_t *out_cnt, int32_t *out_ids);

/* replaces the query text of find_loose_states, analyze_root_cause.py:51-60 (and, with both bounds
 * equal, find_strict_states): intervals with tmin <= tmin_le[i] and tmax > tmax_gt[i]. */
int rca_state_window(rca_engine_t *e, const int32_t *nodes, const int64_t *tmin_le, const int64_t *tmax_gt, int64_t n,
                     int on_device, int32_t *out_first, int32_t *out_cnt, int32_t *out_ids);

/* The same with the label of the state node applied: `(n2:KIND)` of analyze_root_cause.py:72 - only STATE nodes
 * of kind `state_kind` are counted and returned (-1 = any kind), so the cut at RCA_STATE_LIMIT falls after the
 * label test, as the reference's `RETURN n2 LIMIT 10` does (analyze_root_cause.py:76-77). */
int rca_state_window_kind(rca_engine_t *e, const int32_t *nodes, const int64_t *tmin_le, const int64_t *tmax_gt, int64_t n,
                          int on_device, int32_t state_kind, int32_t *out_first, int32_t *out_cnt, int32_t *out_ids);

/* Message arena for rca_text_contains: n_msgs UTF-8 strings, concatenated in `bytes`, message i =
 * bytes[off[i] .. off[i+1]), belonging to node node_ids[i] (ascending node ids give 'lowest id first'
 * answers).  Copied to the device; replaces nothing in the reference - there the text lives in Neo4j. */
int rca_text_load(rca_engine_t *e, const uint8_t *bytes, const int64_t *off, const int32_t *node_ids, int64_t n_msgs);
/* replaces the predicate `N1.message CONTAINS $message` of find_srckind_metapath_neo4j.py:78 and
 * `evt.message CONTAINS <msg>` of generate_query.py:225, for n_pat patterns at once (case-sensitive
 * byte-wise substring test).  out_first_node[i] = lowest node id whose message contains pattern i, or
 * -1 (the canonical answer to 'WITH evt LIMIT 1').  out_pairs (nullable): up to pairs_cap (pattern
 * index, node id) pairs in unspecified order; *n_pairs = matches found, which may exceed pairs_cap
 * (the list is then truncated - call again with a larger buffer).  An empty pattern matches every
 * message; it is reported through out_first_node only. */
int rca_text_contains(rca_engine_t *e, const uint8_t *pat_bytes, const int64_t *pat_off, int64_t n_pat,
                      int32_t *out_first_node, int32_t *out_pairs, int64_t pairs_cap, int64_t *n_pairs);
/* replaces message_compatible, generate_query/generate_query.py:104-129 (the filter run_and_filter_query
 * applies to every row of an instance query: keep the row iff the destination's name, or its kind, occurs in
 * the event's message), for n_pairs rows at once.  Messages are those of the loaded arena (rca_text_load),
 * addressed by their index in load order; needles (names, kind strings) come in one arena like patterns
 * do.  out_flags[k] = 1 iff needle needle_a[k] or needle needle_b[k] occurs in message msg_index[k]
 * (byte-wise, case-sensitive; the empty needle occurs everywhere; index -1 = no such needle). */
int rca_text_match_pairs(rca_engine_t *e, const uint8_t *needle_bytes, const int64_t *needle_off, int64_t n_needles,
                         const int32_t *msg_index, const int32_t *needle_a, const int32_t *needle_b, int64_t n_pairs,
                         uint8_t *out_flags);

/* ---- Count exchange over peer memory -------------------------------------------------------------
 * The one collective of the path (every rank learns every rank's instance count; the reference has no
 * counterpart - it is one process) done by the GPUs themselves over NVLink / NVSwitch: each rank stores its
 * value into a slot of every peer's array and waits until its own array is complete.  One launch of
 * `world` threads on the engine's stream; nothing on the host, no NCCL call on the step's critical path.
 * Ranks are processes of ONE box (at most 8); the arrays are shared with CUDA IPC handles, which the caller
 * carries between the processes (any channel: torch.distributed, a file, a pipe).
 *   rca_peer_handle   allocate this engine's array, return its handle (RCA_PEER_HANDLE_BYTES bytes)
 *   rca_peer_connect  map the arrays of all ranks (handles in rank order, this rank's own included)
 *   rca_peer_allgather publish `value`; out_values != NULL: wait and return all `world` values (rank
 *                     order).  NULL: only enqueue - the value is published and the PREVIOUS exchange is
 *                     waited for on the device, so the ranks stay within one exchange of each other without
 *                     waiting for the slowest one at every step; rca_peer_wait collects the last exchange.
 *                     Collective: every rank calls it the same number of times.  A peer that does not show
 *                     up within the timeout (two seconds unless rca_peer_set_timeout says otherwise) makes the
 *                     call fail instead of hanging the GPU; the error is reported once and the ranks' exchange
 *                     counts may then disagree - set the exchange up again (new engines) or stop.
 *   rca_peer_set_timeout  how long a rank waits for a peer, in milliseconds (first launches, arena growth on
 *                     one rank and similar skew are legitimate reasons to wait longer than the default) */
#define RCA_PEER_HANDLE_BYTES 64
int rca_peer_handle(rca_engine_t *e, int32_t world, uint8_t *out_handle);
int rca_peer_connect(rca_engine_t *e, int32_t rank, const uint8_t *handles);
int rca_peer_allgather(rca_engine_t *e, int64_t value, int64_t *out_values);
int rca_peer_wait(rca_engine_t *e, int64_t *out_values);
int rca_peer_set_timeout(rca_engine_t *e, int64_t milliseconds);

/* ---- Row gather over peer memory ---------------------------------------------------------------------
 * The payload half of the multi-GPU path (the reference has no counterpart - it is one process): after a
 * seed-sharded expansion every rank pushes its rows of one level into a window in the destination rank's
 * device memory, at the offset the exclusive scan of the ranks' counts gives it, so the window holds the rows
 * of all ranks in rank order (= seed order when the shards are contiguous ranges of a sorted seed list).
 * Remote 16-byte stores over NVLink / NVSwitch by a kernel on the engine's stream; counts and completion
 * travel through the count exchange above, which must be set up first (rca_peer_handle / rca_peer_connect).
 *   rca_gather_window   on the destination rank: allocate (or replace) the window, return its CUDA IPC handle
 *   rca_gather_connect  on every rank (the destination included): map the destination's window.  The caller
 *                       carries the handle between the processes, as for rca_peer_connect
 *   rca_gather_rows     collective: gather `level` of the selected plan of the last submission to word
 *                       `window_offset_words` of the window (the same on every rank; several levels can
 *                       share a window).  On the destination *dev_rows points at the rows in the window and
 *                       *n_rows is the total; rank_rows (nullable, [world]) receives every rank's count;
 *                       push_ms (nullable) this rank's device time for its push.  Rows that do not fit the
 *                       window make EVERY rank return RCA_ERR_NOMEM (same totals everywhere): replace the
 *                       window with a larger one and gather again. */
int rca_gather_window(rca_engine_t *e, int64_t words, uint8_t *out_handle);
int rca_gather_connect(rca_engine_t *e, int32_t dst_rank, const uint8_t *handle, int64_t words);
int rca_gather_rows(rca_engine_t *e, int level, int64_t window_offset_words, const int32_t **dev_rows, int64_t *n_rows,
                    int64_t *rank_rows, double *push_ms);

/* Limits of rca_expand_static: the fixed scratch holds at most this many seeds and this many
 * partial paths (summed over all hops of one call), in RCA_STATIC_SCRATCH_BYTES of device memory
 * that every GPU which loads the library keeps until the process exits. */
#define RCA_STATIC_MAX_SEEDS 262144
#define RCA_STATIC_MAX_FRONTIER 262144
#define RCA_STATIC_SCRATCH_BYTES (5u << 20)

/* No counterpart in the reference.  rca_expand for callers that cannot allocate device memory:
 * a service whose device-memory budget was handed to a pool at start-up, a latency path where
 * cudaMalloc (which may synchronise the whole device) is not allowed, a process under a memory
 * cap.  Same plans and the same set of rows as rca_expand, but
 *   - it calls no cudaMalloc / cudaFree and no allocating library routine: its scratch is a fixed
 *     array inside the library, limited by the three constants above; a plan that needs more is
 *     refused with RCA_ERR_ARG / RCA_ERR_NOMEM and a message naming the limit - nothing grows;
 *   - seeds, timestamps and rows are the CALLER's DEVICE buffers: dev_seeds[n_seeds],
 *     dev_seed_ts[n_seeds] (only with plan->with_states), dev_rows[rows_capacity_words].  Rows of
 *     length 1 come first, then length 2, ...; stats->emitted[s] and stats->row_len[s] say how
 *     many and how wide, so level s starts at word sum_{j<s} emitted[j] * row_len[j].  Within a
 *     level the rows are in the order of the single-launch-per-hop kernels, i.e. the order
 *     rca_expand gives below its fused threshold (rca_set_fused_threshold); above that threshold
 *     rca_expand may order a level differently - same rows, compare as sets;
 *   - the kernels run on the ENGINE's stream (its own non-blocking stream unless rca_set_stream
 *     gave it another), which is not ordered against the stream that produced the caller's
 *     buffers: either hand the engine that stream with rca_set_stream, or finish the work that
 *     writes dev_seeds / dev_seed_ts / dev_rows (synchronise, or an event wait) before the call.
 *     The call returns after it has waited for the engine's stream and read the counters back, so
 *     the rows are complete when it returns;
 *   - the first call on a device looks up the scratch array's device address; make that first
 *     call before starting any stream capture;
 *   - all calls on one GPU share the one scratch array, which keeps what the previous call left in
 *     it (the call resets what it needs).  Two calls that overlap on the same GPU - from two
 *     engines, threads or streams - corrupt each other: run them one after the other. */
int rca_expand_static(rca_engine_t *e, const rca_plan_t *plan, const int32_t *dev_seeds, int64_t n_seeds,
                      const int64_t *dev_seed_ts, int32_t *dev_rows, int64_t rows_capacity_words,
                      rca_stats_t *stats);

/* library / build information, e.g. "rca_b200 0.1 sm_100a" */
const char *rca_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RCA_B200_H */
