/* CPU ORACLE - TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement, over the snapshot's CSR arrays, of the same query semantics that
 * oracle/cypher_oracle.py states over a property graph (which in turn restates the Cypher the
 * reference sends to Neo4j: find_metapath/find_srckind_metapath_neo4j.py:95-129,
 * generate_query/generate_query.py:241-260, check_state/analyze_root_cause.py:70-79).
 * It exists so that parity can be checked at sizes the Python oracle cannot finish, and so that
 * bench.py has a CPU baseline (kind "port") to time on the GPU box's host cores.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.  PARITY UNPINNED - see the header of cypher_oracle.py; this file is pinned to
 * cypher_oracle.py by tests/test_oracle.py.
 *
 * Deliberately independent of the engine's derived tables: undirected hops are composed from the
 * out-CSR and the in-CSC (not the undirected table), typed hops filter full rows by rel_et
 * (not the typed-range tables), and there is no reachability pruning - a depth-first walk per
 * seed, exactly the pruned form in cypher_oracle.py::varlen_paths.
 *
 * Row format (both functions): per path of length s, (s+1) node ids then s relationship ids,
 * appended to the buffer of level s.  Row order is unspecified; callers sort.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_MAX_HOPS 8
#define ORC_REL_MASK 0x3fffffff   /* snapshot format: bit 30 of an in_adj rel word marks a reversed hop */

typedef struct {
    int32_t n_nodes, n_kinds;
    const int32_t *kind_start;          /* [n_kinds+1] */
    const int32_t *out_rng;             /* [n_nodes+1][2] (begin, entity_end) */
    const int32_t *out_dst;             /* [E] */
    const int32_t *in_rng;              /* [n_nodes+1][2] */
    const int32_t *in_adj;              /* [E][2] (src, rel) */
    const int32_t *rel_et;              /* [E] */
    const int32_t *st_off;              /* [n_nodes+1] */
    const int64_t *st_tmin, *st_tmax;   /* [S] */
    const int32_t *st_node;             /* [S] */
} orc_graph_t;

typedef struct {
    int32_t direction;                  /* 0 out, 1 in, 2 undirected */
    int32_t min_hops, max_hops;
    int32_t entity_only;                /* walk only the entity-entity part of each row */
    int32_t node_simple;                /* no node repeats */
    int32_t dest_kind;                  /* -1: any */
    int32_t has_inter;                  /* apply the any(interior in inter) rule */
    const uint8_t *excl;                /* [n_kinds] 1 = kind may not appear anywhere on the path */
    const uint8_t *inter;               /* [n_kinds] */
    const int32_t *pos_kind;            /* [max_hops+1] kind required at position p, -1 any; may be NULL */
} orc_varlen_t;

typedef struct { int32_t *p; int64_t n, cap; } buf_t;

static void buf_push(buf_t *b, const int32_t *nodes, const int32_t *rels, int s) {
    int64_t need = b->n + 2 * s + 1;
    if (need > b->cap) {
        b->cap = need * 2 + 1024;
        b->p = (int32_t *)realloc(b->p, (size_t)b->cap * sizeof(int32_t));
    }
    memcpy(b->p + b->n, nodes, (size_t)(s + 1) * sizeof(int32_t));
    memcpy(b->p + b->n + s + 1, rels, (size_t)s * sizeof(int32_t));
    b->n = need;
}

static int kind_of(const orc_graph_t *g, int32_t v) {
    int lo = 0, hi = g->n_kinds;        /* kind_start[lo] <= v < kind_start[hi] */
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (g->kind_start[mid] <= v) lo = mid; else hi = mid;
    }
    return lo;
}

typedef struct {
    const orc_graph_t *g;
    const orc_varlen_t *q;
    buf_t *lv;                          /* [ORC_MAX_HOPS+1] */
    int64_t scanned;
    int32_t nodes[ORC_MAX_HOPS + 1], rels[ORC_MAX_HOPS];
    int32_t kinds[ORC_MAX_HOPS + 1];
} walk_t;

static void varlen_visit(walk_t *w, int s);

static void varlen_try(walk_t *w, int s, int32_t v, int32_t r) {
    const orc_varlen_t *q = w->q;
    w->scanned++;
    for (int i = 0; i < s; i++) if (w->rels[i] == r) return;          /* relationship uniqueness */
    if (q->node_simple) for (int i = 0; i <= s; i++) if (w->nodes[i] == v) return;
    int k = kind_of(w->g, v);
    if (q->excl && q->excl[k]) return;
    if (q->pos_kind && q->pos_kind[s + 1] >= 0 && q->pos_kind[s + 1] != k) return;
    w->nodes[s + 1] = v; w->rels[s] = r; w->kinds[s + 1] = k;
    varlen_visit(w, s + 1);
}

static void varlen_visit(walk_t *w, int s) {
    const orc_graph_t *g = w->g;
    const orc_varlen_t *q = w->q;
    if (s >= q->min_hops && (q->dest_kind < 0 || w->kinds[s] == q->dest_kind)) {
        int ok = 1;
        if (q->has_inter) {
            ok = 0;
            for (int i = 1; i < s; i++) if (q->inter[w->kinds[i]]) { ok = 1; break; }
        }
        if (ok) buf_push(&w->lv[s], w->nodes, w->rels, s);
    }
    if (s == q->max_hops) return;
    int32_t u = w->nodes[s];
    if (q->direction == 0 || q->direction == 2) {
        int32_t b = g->out_rng[2 * u], e = q->entity_only ? g->out_rng[2 * u + 1] : g->out_rng[2 * (u + 1)];
        for (int32_t j = b; j < e; j++) varlen_try(w, s, g->out_dst[j], j);
    }
    if (q->direction == 1 || q->direction == 2) {
        int32_t b = g->in_rng[2 * u], e = q->entity_only ? g->in_rng[2 * u + 1] : g->in_rng[2 * (u + 1)];
        for (int32_t j = b; j < e; j++) {
            int32_t v = g->in_adj[2 * j], r = g->in_adj[2 * j + 1] & ORC_REL_MASK;
            if (q->direction == 2 && v == u) continue;                /* a loop is traversed once */
            varlen_try(w, s, v, r);
        }
    }
}

static int64_t gather(buf_t *bufs, int nthreads, int64_t *level_rows, int32_t **out, int max_hops) {
    int64_t total = 0;
    for (int s = 0; s <= max_hops; s++) {
        int64_t n = 0;
        for (int t = 0; t < nthreads; t++) n += bufs[t * (ORC_MAX_HOPS + 1) + s].n;
        level_rows[s] = n / (2 * s + 1);
        total += level_rows[s];
        out[s] = n ? (int32_t *)malloc((size_t)n * sizeof(int32_t)) : NULL;
        int64_t off = 0;
        for (int t = 0; t < nthreads; t++) {
            buf_t *b = &bufs[t * (ORC_MAX_HOPS + 1) + s];
            if (b->n) memcpy(out[s] + off, b->p, (size_t)b->n * sizeof(int32_t));
            off += b->n;
            free(b->p);
        }
    }
    free(bufs);
    return total;
}

/* Variable-length enumeration from explicit seeds.  level_rows[s], out[s] (malloc'ed, free with
 * orc_free) for s = 0..max_hops.  Returns the total number of paths, or -1 on bad arguments. */
int64_t orc_varlen(const orc_graph_t *g, const orc_varlen_t *q, const int32_t *seeds, int64_t n_seeds,
                   int threads, int64_t *level_rows, int32_t **out, int64_t *scanned) {
    if (q->max_hops > ORC_MAX_HOPS || q->min_hops < 0) return -1;
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#else
    threads = 1;
#endif
    buf_t *bufs = (buf_t *)calloc((size_t)threads * (ORC_MAX_HOPS + 1), sizeof(buf_t));
    int64_t sc = 0;
#pragma omp parallel num_threads(threads) reduction(+ : sc)
    {
#ifdef _OPENMP
        int tid = omp_get_thread_num();
#else
        int tid = 0;
#endif
        walk_t w;
        w.g = g; w.q = q; w.lv = bufs + (size_t)tid * (ORC_MAX_HOPS + 1); w.scanned = 0;
#pragma omp for schedule(dynamic, 64)
        for (int64_t i = 0; i < n_seeds; i++) {
            int32_t v = seeds[i];
            int k = kind_of(g, v);
            if (q->excl && q->excl[k]) continue;
            if (q->pos_kind && q->pos_kind[0] >= 0 && q->pos_kind[0] != k) continue;
            w.nodes[0] = v; w.kinds[0] = k;
            varlen_visit(&w, 0);
        }
        sc += w.scanned;
    }
    if (scanned) *scanned = sc;
    return gather(bufs, threads, level_rows, out, q->max_hops);
}

/* Typed chain from explicit seeds: step i follows relationships of edge type et[i], forward
 * (dir[i] = 0: current node is the start) or backward (dir[i] = 1: current node is the end).
 * No uniqueness between steps (each step is its own MATCH clause).  One level of output: L. */
int64_t orc_chain(const orc_graph_t *g, const int32_t *et, const int32_t *dir, int n_steps,
                  const int32_t *seeds, int64_t n_seeds, int threads, int32_t **out, int64_t *scanned) {
    if (n_steps > ORC_MAX_HOPS || n_steps < 0) return -1;
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#else
    threads = 1;
#endif
    buf_t *bufs = (buf_t *)calloc((size_t)threads * (ORC_MAX_HOPS + 1), sizeof(buf_t));
    int64_t sc = 0;
#pragma omp parallel num_threads(threads) reduction(+ : sc)
    {
#ifdef _OPENMP
        int tid = omp_get_thread_num();
#else
        int tid = 0;
#endif
        buf_t *lv = bufs + (size_t)tid * (ORC_MAX_HOPS + 1);
        int32_t nodes[ORC_MAX_HOPS + 1], rels[ORC_MAX_HOPS], cur[ORC_MAX_HOPS], end[ORC_MAX_HOPS];
        int64_t mysc = 0;
#pragma omp for schedule(dynamic, 256)
        for (int64_t i = 0; i < n_seeds; i++) {
            nodes[0] = seeds[i];
            if (n_steps == 0) { buf_push(&lv[0], nodes, rels, 0); continue; }
            int s = 0;
            const int32_t *rng = dir[0] ? g->in_rng : g->out_rng;
            cur[0] = rng[2 * nodes[0]]; end[0] = rng[2 * (nodes[0] + 1)];
            while (s >= 0) {
                if (cur[s] >= end[s]) { s--; continue; }
                int32_t j = cur[s]++, v, r;
                mysc++;
                if (dir[s]) { v = g->in_adj[2 * j]; r = g->in_adj[2 * j + 1] & ORC_REL_MASK; } else { v = g->out_dst[j]; r = j; }
                if (g->rel_et[r] != et[s]) continue;
                nodes[s + 1] = v; rels[s] = r;
                if (s + 1 == n_steps) { buf_push(&lv[n_steps], nodes, rels, n_steps); continue; }
                s++;
                rng = dir[s] ? g->in_rng : g->out_rng;
                cur[s] = rng[2 * v]; end[s] = rng[2 * (v + 1)];
            }
        }
        sc += mysc;
    }
    if (scanned) *scanned = sc;
    int64_t level_rows[ORC_MAX_HOPS + 1];
    int32_t *outs[ORC_MAX_HOPS + 1];
    gather(bufs, threads, level_rows, outs, n_steps);
    for (int s = 0; s < n_steps; s++) free(outs[s]);
    *out = outs[n_steps];
    return level_rows[n_steps];
}

/* HasState predicate, analyze_root_cause.py:70-79: for query i, intervals of entity nodes[i]
 * with tmin <= ts[i] and tmax > ts_hi[i] (ts_hi == ts: the strict form; :51-60 is the loose form).  first[i] = lowest matching STATE node id or -1, cnt[i] = number
 * of matches; ids (may be NULL) receives up to `limit` ids per query, -1 padded. */
void orc_states(const orc_graph_t *g, const int32_t *nodes, const int64_t *ts, const int64_t *ts_hi, int64_t n, int limit,
                int32_t *first, int32_t *cnt, int32_t *ids, int threads) {
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#else
    threads = 1;
#endif
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t i = 0; i < n; i++) {
        int32_t v = nodes[i], c = 0, f = -1;
        for (int32_t j = g->st_off[v]; j < g->st_off[v + 1]; j++) {
            if (g->st_tmin[j] <= ts[i] && g->st_tmax[j] > ts_hi[i]) {
                if (c == 0) f = g->st_node[j];
                if (ids && c < limit) ids[i * limit + c] = g->st_node[j];
                c++;
            }
        }
        if (ids) for (int32_t k = c; k < limit; k++) ids[i * limit + k] = -1;
        first[i] = f; cnt[i] = c;
    }
}

/* ================================================================================================
 * Whole-step variants for the CPU baseline (bench.py: cpu_baseline / --impl reference).
 *
 * Same semantics as orc_varlen / orc_chain above, with what a careful CPU implementation of the SAME
 * job would also do, so that the timed figure is not a straw man:
 *   - the state predicate (analyze_root_cause.py:70-79) is evaluated here, per node of every row,
 *     against the seed's own timestamp - rows come out as (s+1) nodes, s relationships, (s+1) first
 *     state ids, 1 missing-state mask: the row format of the engine's with_states plans - and a
 *     node's answer is kept while siblings share that node (once per partial path, not per row);
 *   - node -> kind is one byte load from a table built once per call (not a binary search);
 *   - optional schema pruning (`reach` non-NULL): a partial path is dropped when its tail kind cannot
 *     reach dest_kind in the hops that remain.  Result-neutral; it is the kind-level part of what the
 *     engine's plans do (rca_plan_t::prune).  With reach == NULL nothing is pruned, as in orc_varlen.
 * tests/test_oracle.py holds both variants to orc_varlen / orc_chain + orc_states.
 */
typedef struct {
    const orc_graph_t *g;
    const orc_varlen_t *q;
    buf_t *lv;
    int64_t scanned;
    const uint8_t *nk;                  /* node -> kind */
    const uint32_t *reach;              /* [n_kinds] bit r-1: dest_kind is reachable in 1..r hops; NULL: no pruning */
    int64_t ts;
    int with_states;
    int32_t nodes[ORC_MAX_HOPS + 1], rels[ORC_MAX_HOPS], kinds[ORC_MAX_HOPS + 1];
    int32_t st_first[ORC_MAX_HOPS + 1];
    uint8_t st_known[ORC_MAX_HOPS + 1];
} walk2_t;

static int32_t state_first(const orc_graph_t *g, int32_t v, int64_t ts) {
    for (int32_t j = g->st_off[v]; j < g->st_off[v + 1]; j++)
        if (g->st_tmin[j] <= ts && g->st_tmax[j] > ts) return g->st_node[j];      /* records are sorted by STATE node id */
    return -1;
}

static void emit2(walk2_t *w, int s) {
    const int width = w->with_states ? 3 * s + 3 : 2 * s + 1;
    buf_t *b = &w->lv[s];
    int64_t need = b->n + width;
    if (need > b->cap) {
        b->cap = need * 2 + 4096;
        b->p = (int32_t *)realloc(b->p, (size_t)b->cap * sizeof(int32_t));
    }
    int32_t *row = b->p + b->n;
    memcpy(row, w->nodes, (size_t)(s + 1) * sizeof(int32_t));
    memcpy(row + s + 1, w->rels, (size_t)s * sizeof(int32_t));
    if (w->with_states) {
        uint32_t miss = 0;
        for (int d = 0; d <= s; d++) {
            if (!w->st_known[d]) { w->st_first[d] = state_first(w->g, w->nodes[d], w->ts); w->st_known[d] = 1; }
            row[2 * s + 1 + d] = w->st_first[d];
            if (w->st_first[d] < 0) miss |= 1u << d;
        }
        row[3 * s + 2] = (int32_t)miss;
    }
    b->n = need;
}

static void visit2(walk2_t *w, int s);

static inline void try2(walk2_t *w, int s, int32_t v, int32_t r) {
    const orc_varlen_t *q = w->q;
    w->scanned++;
    for (int i = 0; i < s; i++) if (w->rels[i] == r) return;
    if (q->node_simple) for (int i = 0; i <= s; i++) if (w->nodes[i] == v) return;
    const int k = w->nk[v];
    if (q->excl && q->excl[k]) return;
    if (q->pos_kind && q->pos_kind[s + 1] >= 0 && q->pos_kind[s + 1] != k) return;
    if (w->reach) {                     /* neither a result here nor on the way to one */
        const int left = q->max_hops - (s + 1);
        const int emits = s + 1 >= q->min_hops && k == q->dest_kind;
        const int goes_on = left > 0 && ((w->reach[k] >> (left - 1)) & 1u);
        if (!emits && !goes_on) return;
    }
    w->nodes[s + 1] = v; w->rels[s] = r; w->kinds[s + 1] = k; w->st_known[s + 1] = 0;
    visit2(w, s + 1);
}

static void visit2(walk2_t *w, int s) {
    const orc_graph_t *g = w->g;
    const orc_varlen_t *q = w->q;
    if (s >= q->min_hops && (q->dest_kind < 0 || w->kinds[s] == q->dest_kind)) {
        int ok = 1;
        if (q->has_inter) {
            ok = 0;
            for (int i = 1; i < s; i++) if (q->inter[w->kinds[i]]) { ok = 1; break; }
        }
        if (ok) emit2(w, s);
    }
    if (s == q->max_hops) return;
    if (w->reach && !((w->reach[w->kinds[s]] >> (q->max_hops - s - 1)) & 1u)) return;
    const int32_t u = w->nodes[s];
    if (q->direction == 0 || q->direction == 2) {
        const int32_t b = g->out_rng[2 * u], e = q->entity_only ? g->out_rng[2 * u + 1] : g->out_rng[2 * (u + 1)];
        for (int32_t j = b; j < e; j++) try2(w, s, g->out_dst[j], j);
    }
    if (q->direction == 1 || q->direction == 2) {
        const int32_t b = g->in_rng[2 * u], e = q->entity_only ? g->in_rng[2 * u + 1] : g->in_rng[2 * (u + 1)];
        for (int32_t j = b; j < e; j++) {
            const int32_t v = g->in_adj[2 * j], r = g->in_adj[2 * j + 1] & ORC_REL_MASK;
            if (q->direction == 2 && v == u) continue;
            try2(w, s, v, r);
        }
    }
}

static uint8_t *node_kinds(const orc_graph_t *g) {
    uint8_t *nk = (uint8_t *)malloc((size_t)g->n_nodes + 1);
    for (int k = 0; k < g->n_kinds; k++)
        memset(nk + g->kind_start[k], k, (size_t)(g->kind_start[k + 1] - g->kind_start[k]));
    return nk;
}

/* Kind-level reachability of dest_kind, measured from the data: reach[k] bit r-1 is set when some walk of 1..r hops
 * in the query's direction, through kinds the query does not exclude, leads from kind k to dest_kind.  Returns 0 when
 * pruning does not apply (no destination, positional kind constraints, more than 255 kinds). */
int orc_reach(const orc_graph_t *g, const orc_varlen_t *q, uint32_t *reach) {
    const int nk = g->n_kinds;
    if (q->dest_kind < 0 || nk > 255) return 0;
    if (q->pos_kind) for (int i = 1; i <= q->max_hops; i++) if (q->pos_kind[i] >= 0) return 0;
    uint8_t *kind = node_kinds(g);
    uint8_t *adj = (uint8_t *)calloc((size_t)nk * nk, 1);       /* adj[a*nk+b]: a hop a -> b exists in walk direction */
    for (int32_t u = 0; u < g->n_nodes; u++) {
        const int32_t b = g->out_rng[2 * u], e = q->entity_only ? g->out_rng[2 * u + 1] : g->out_rng[2 * (u + 1)];
        for (int32_t j = b; j < e; j++) {
            const int a = kind[u], c = kind[g->out_dst[j]];
            if (q->direction != 1) adj[a * nk + c] = 1;
            if (q->direction != 0) adj[c * nk + a] = 1;
        }
    }
    uint8_t *can = (uint8_t *)calloc((size_t)nk, 1), *nxt = (uint8_t *)calloc((size_t)nk, 1);
    for (int k = 0; k < nk; k++) reach[k] = 0;
    if (!(q->excl && q->excl[q->dest_kind])) {
        for (int r = 1; r <= q->max_hops; r++) {
            memset(nxt, 0, (size_t)nk);
            for (int a = 0; a < nk; a++)
                for (int c = 0; c < nk; c++) {
                    if (!adj[a * nk + c]) continue;
                    if (c == q->dest_kind || (r > 1 && can[c] && !(q->excl && q->excl[c]))) nxt[a] = 1;
                }
            for (int k = 0; k < nk; k++) {
                can[k] |= nxt[k];
                if (can[k]) reach[k] |= 1u << (r - 1);
            }
        }
    }
    free(kind); free(adj); free(can); free(nxt);
    return 1;
}

/* orc_varlen with per-seed timestamps (seed_ts non-NULL: rows carry the state columns) and optional pruning. */
int64_t orc_varlen_step(const orc_graph_t *g, const orc_varlen_t *q, const int32_t *seeds, const int64_t *seed_ts,
                        int64_t n_seeds, const uint32_t *reach, int threads, int64_t *level_rows, int32_t **out,
                        int64_t *scanned) {
    if (q->max_hops > ORC_MAX_HOPS || q->min_hops < 0) return -1;
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#else
    threads = 1;
#endif
    uint8_t *nk = node_kinds(g);
    buf_t *bufs = (buf_t *)calloc((size_t)threads * (ORC_MAX_HOPS + 1), sizeof(buf_t));
    int64_t sc = 0;
#pragma omp parallel num_threads(threads) reduction(+ : sc)
    {
#ifdef _OPENMP
        int tid = omp_get_thread_num();
#else
        int tid = 0;
#endif
        walk2_t w;
        memset(&w, 0, sizeof w);
        w.g = g; w.q = q; w.lv = bufs + (size_t)tid * (ORC_MAX_HOPS + 1); w.nk = nk; w.reach = reach;
        w.with_states = seed_ts != NULL;
#pragma omp for schedule(dynamic, 64)
        for (int64_t i = 0; i < n_seeds; i++) {
            const int32_t v = seeds[i];
            if (v < 0 || v >= g->n_nodes) continue;
            const int k = nk[v];
            if (q->excl && q->excl[k]) continue;
            if (q->pos_kind && q->pos_kind[0] >= 0 && q->pos_kind[0] != k) continue;
            w.nodes[0] = v; w.kinds[0] = k; w.st_known[0] = 0;
            w.ts = seed_ts ? seed_ts[i] : 0;
            visit2(&w, 0);
        }
        sc += w.scanned;
    }
    free(nk);
    if (scanned) *scanned = sc;
    /* gather(): its row width is 2s+1; here rows may be 3s+3 wide */
    int64_t total = 0;
    for (int s = 0; s <= q->max_hops; s++) {
        const int width = seed_ts ? 3 * s + 3 : 2 * s + 1;
        int64_t n = 0;
        for (int t = 0; t < threads; t++) n += bufs[t * (ORC_MAX_HOPS + 1) + s].n;
        level_rows[s] = n / width;
        total += level_rows[s];
        out[s] = n ? (int32_t *)malloc((size_t)n * sizeof(int32_t)) : NULL;
        int64_t off = 0;
        for (int t = 0; t < threads; t++) {
            buf_t *b = &bufs[t * (ORC_MAX_HOPS + 1) + s];
            if (b->n) memcpy(out[s] + off, b->p, (size_t)b->n * sizeof(int32_t));
            off += b->n;
            free(b->p);
        }
    }
    free(bufs);
    return total;
}

/* orc_chain with per-seed timestamps: rows carry the state columns when seed_ts is non-NULL.  The walk itself is
 * orc_chain's (full rows filtered by edge type - the oracle keeps no typed tables). */
int64_t orc_chain_step(const orc_graph_t *g, const int32_t *et, const int32_t *dir, int n_steps, const int32_t *seeds,
                       const int64_t *seed_ts, int64_t n_seeds, int threads, int32_t **out, int64_t *scanned) {
    if (n_steps > ORC_MAX_HOPS || n_steps < 1) return -1;
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#else
    threads = 1;
#endif
    const int L = n_steps, width = seed_ts ? 3 * L + 3 : 2 * L + 1;
    buf_t *bufs = (buf_t *)calloc((size_t)threads, sizeof(buf_t));
    int64_t sc = 0;
#pragma omp parallel num_threads(threads) reduction(+ : sc)
    {
#ifdef _OPENMP
        int tid = omp_get_thread_num();
#else
        int tid = 0;
#endif
        buf_t *b = bufs + tid;
        int32_t nodes[ORC_MAX_HOPS + 1], rels[ORC_MAX_HOPS], cur[ORC_MAX_HOPS], end[ORC_MAX_HOPS], stf[ORC_MAX_HOPS + 1];
        uint8_t known[ORC_MAX_HOPS + 1];
        int64_t mysc = 0;
#pragma omp for schedule(dynamic, 256)
        for (int64_t i = 0; i < n_seeds; i++) {
            nodes[0] = seeds[i];
            if (nodes[0] < 0 || nodes[0] >= g->n_nodes) continue;
            known[0] = 0;
            const int64_t ts = seed_ts ? seed_ts[i] : 0;
            int s = 0;
            const int32_t *rng = dir[0] ? g->in_rng : g->out_rng;
            cur[0] = rng[2 * nodes[0]]; end[0] = rng[2 * (nodes[0] + 1)];
            while (s >= 0) {
                if (cur[s] >= end[s]) { s--; continue; }
                int32_t j = cur[s]++, v, r;
                mysc++;
                if (dir[s]) { v = g->in_adj[2 * j]; r = g->in_adj[2 * j + 1] & ORC_REL_MASK; } else { v = g->out_dst[j]; r = j; }
                if (g->rel_et[r] != et[s]) continue;
                nodes[s + 1] = v; rels[s] = r; known[s + 1] = 0;
                if (s + 1 == L) {
                    int64_t need = b->n + width;
                    if (need > b->cap) { b->cap = need * 2 + 4096; b->p = (int32_t *)realloc(b->p, (size_t)b->cap * sizeof(int32_t)); }
                    int32_t *row = b->p + b->n;
                    memcpy(row, nodes, (size_t)(L + 1) * sizeof(int32_t));
                    memcpy(row + L + 1, rels, (size_t)L * sizeof(int32_t));
                    if (seed_ts) {
                        uint32_t miss = 0;
                        for (int d = 0; d <= L; d++) {
                            if (!known[d]) { stf[d] = state_first(g, nodes[d], ts); known[d] = 1; }
                            row[2 * L + 1 + d] = stf[d];
                            if (stf[d] < 0) miss |= 1u << d;
                        }
                        row[3 * L + 2] = (int32_t)miss;
                    }
                    b->n = need;
                    continue;
                }
                s++;
                rng = dir[s] ? g->in_rng : g->out_rng;
                cur[s] = rng[2 * v]; end[s] = rng[2 * (v + 1)];
            }
        }
        sc += mysc;
    }
    if (scanned) *scanned = sc;
    int64_t n = 0;
    for (int t = 0; t < threads; t++) n += bufs[t].n;
    *out = n ? (int32_t *)malloc((size_t)n * sizeof(int32_t)) : NULL;
    int64_t off = 0;
    for (int t = 0; t < threads; t++) {
        if (bufs[t].n) memcpy(*out + off, bufs[t].p, (size_t)bufs[t].n * sizeof(int32_t));
        off += bufs[t].n;
        free(bufs[t].p);
    }
    free(bufs);
    return n / width;
}

void orc_free(void *p) { free(p); }

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
