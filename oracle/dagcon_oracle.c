/* TEST INFRASTRUCTURE -- CPU oracle for the DAGCon consensus hot path.
 *
 * A plain-C restatement of the reference algorithm (verdurin/pbdagcon @ 3c382f2):
 *   normalizeGaps  src/cpp/Alignment.cpp:131-217
 *   trimAln        src/cpp/Alignment.cpp:219-242
 *   AlnGraphBoost  src/cpp/AlnGraphBoost.cpp:16-459  (ctor, addAln, addEdge, mergeNodes,
 *                  mergeInNodes, mergeOutNodes, markForReaper, bestPath, consensus x2)
 *   Consensus functor (filters + per-target driver)  src/cpp/main.cpp:111-151, dazcon.cpp:61-107
 *
 * It deliberately keeps the reference's data-structure semantics (per-vertex ordered in/out
 * edge vectors, std::map<char,...> grouping order, FIFO queues, float scores) rather than the
 * product's flattened arena, so that it is an independent check of the CUDA path.
 *
 * PARITY PINNED: tests/test_oracle_*.py check this file against (a) every KAT of the
 * reference's own unit tests (test/cpp/AlignmentTest.cpp, AlnGraphBoostTest.cpp) and
 * (b) oracle/_ref/libdagcon_ref.so = the UNMODIFIED reference sources compiled against
 * oracle/shim, on thousands of fuzzed targets, plus committed golden vectors in tests/golden/.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this.  The product library never links or calls it.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

/* ------------------------------------------------------------------ small containers */
typedef struct { int* v; int n, cap; } ivec;
static void iv_push(ivec* a, int x) {
    if (a->n == a->cap) { a->cap = a->cap ? 2 * a->cap : 4; a->v = (int*)realloc(a->v, sizeof(int) * (size_t)a->cap); }
    a->v[a->n++] = x;
}
/* order-preserving erase of the first occurrence (edge ids are unique in a list) */
static void iv_erase(ivec* a, int x) {
    int i, j;
    for (i = 0; i < a->n; i++) if (a->v[i] == x) break;
    if (i == a->n) return;
    for (j = i + 1; j < a->n; j++) a->v[j - 1] = a->v[j];
    a->n--;
}
static void iv_free(ivec* a) { free(a->v); a->v = 0; a->n = a->cap = 0; }

typedef struct { char* p; int64_t n, cap; } cbuf;
static void cb_putn(cbuf* b, const char* s, int64_t n) {
    if (b->n + n + 1 > b->cap) {
        while (b->n + n + 1 > b->cap) b->cap = b->cap ? 2 * b->cap : 256;
        b->p = (char*)realloc(b->p, (size_t)b->cap);
    }
    memcpy(b->p + b->n, s, (size_t)n); b->n += n; b->p[b->n] = 0;
}
static void cb_putc(cbuf* b, char c) { cb_putn(b, &c, 1); }

/* ------------------------------------------------------------------ normalizeGaps / trimAln */

/* Alignment.cpp:131-217.  qout/tout need room for 2*len bytes.  Returns the new length.
 * len == 0 makes the reference's `len-1` underflow (undefined behaviour, never reached in the
 * pipeline because of the minLen filter); here it yields an empty alignment. */
int64_t orc_normalize_gaps(const char* q, const char* t, int64_t len, char* qout, char* tout) {
    int64_t i, n = 0, m = 0;
    char* qn = (char*)malloc((size_t)(2 * len + 1));
    char* tn = (char*)malloc((size_t)(2 * len + 1));
    /* dots to dashes (:141-144) fused with mismatch -> deletion-then-insertion (:147-159) */
    for (i = 0; i < len; i++) {
        char qb = q[i] == '.' ? '-' : q[i];
        char tb = t[i] == '.' ? '-' : t[i];
        if (qb != tb && qb != '-' && tb != '-') {
            qn[n] = '-'; tn[n] = tb; n++;
            qn[n] = qb;  tn[n] = '-'; n++;
        } else {
            qn[n] = qb; tn[n] = tb; n++;
        }
    }
    /* push gaps to the right, but not past the end (:165-198) */
    if (n > 0) {
        for (i = 0; i < n - 1; i++) {
            if (tn[i] == '-') {
                int64_t j = i;
                while (++j < n) {
                    char c = tn[j];
                    if (c != '-') {
                        if (c == qn[i]) { tn[i] = c; tn[j] = '-'; }
                        break;
                    }
                }
            }
            if (qn[i] == '-') {
                int64_t j = i;
                while (++j < n) {
                    char c = qn[j];
                    if (c != '-') {
                        if (c == tn[i]) { qn[i] = c; qn[j] = '-'; }
                        break;
                    }
                }
            }
        }
    }
    /* drop all-gap columns (:210-215) */
    for (i = 0; i < n; i++) {
        if (qn[i] != '-' || tn[i] != '-') { qout[m] = qn[i]; tout[m] = tn[i]; m++; }
    }
    free(qn); free(tn);
    return m;
}

/* Alignment.cpp:219-242.  Returns the new length, *loffs_out = left offset of the kept
 * window in the input strings, *start_io += number of target bases trimmed on the left. */
int64_t orc_trim(const char* t, int64_t len, int trim_len, uint32_t* start_io, int64_t* loffs_out) {
    int lbases = 0, rbases = 0;
    int64_t loffs = 0, roffs = len;
    while (lbases < trim_len && loffs < len) {
        if (t[loffs++] != '-') lbases++;
    }
    while (rbases < trim_len && roffs > loffs) {
        if (t[--roffs] != '-') rbases++;
    }
    *start_io += (uint32_t)lbases;
    *loffs_out = loffs;
    return roffs - loffs;
}

/* ------------------------------------------------------------------ the alignment graph */
typedef struct {
    char base; int coverage; int weight; char backbone; char deleted;
    int bbmap;          /* _bbMap[v]; absent keys read as 0 (std::map::operator[]) */
    ivec out, in;       /* edge ids, insertion order, order-preserving erase */
} onode;
typedef struct { int src, dst, count; char visited; char alive; } oedge;
typedef struct {
    onode* nodes; int nv, capv;
    oedge* edges; int ne, cape;
    int enter, exit_;
} ograph;

static int g_add_vertex(ograph* g) {
    onode* n;
    if (g->nv == g->capv) {
        g->capv = g->capv ? 2 * g->capv : 64;
        g->nodes = (onode*)realloc(g->nodes, sizeof(onode) * (size_t)g->capv);
    }
    n = &g->nodes[g->nv];
    memset(n, 0, sizeof *n);
    n->base = 'N';                      /* AlnNode() defaults, AlnGraphBoost.hpp:30-36 */
    return g->nv++;
}
static int g_add_edge(ograph* g, int u, int v) {
    oedge* e;
    if (g->ne == g->cape) {
        g->cape = g->cape ? 2 * g->cape : 64;
        g->edges = (oedge*)realloc(g->edges, sizeof(oedge) * (size_t)g->cape);
    }
    e = &g->edges[g->ne];
    e->src = u; e->dst = v; e->count = 0; e->visited = 0; e->alive = 1;   /* AlnEdge() */
    iv_push(&g->nodes[u].out, g->ne);
    iv_push(&g->nodes[v].in, g->ne);
    return g->ne++;
}
/* boost::edge(u,v,g): first match in out(u) */
static int g_find_edge(ograph* g, int u, int v) {
    int i;
    for (i = 0; i < g->nodes[u].out.n; i++)
        if (g->edges[g->nodes[u].out.v[i]].dst == v) return g->nodes[u].out.v[i];
    return -1;
}
/* boost::clear_vertex on a bidirectional graph */
static void g_clear_vertex(ograph* g, int u) {
    int i;
    onode* n = &g->nodes[u];
    for (i = 0; i < n->out.n; i++) {
        int e = n->out.v[i];
        if (g->edges[e].dst != u) iv_erase(&g->nodes[g->edges[e].dst].in, e);
        g->edges[e].alive = 0;
    }
    for (i = 0; i < n->in.n; i++) {
        int e = n->in.v[i];
        if (g->edges[e].src != u) iv_erase(&g->nodes[g->edges[e].src].out, e);
        g->edges[e].alive = 0;
    }
    n->out.n = 0; n->in.n = 0;
}
static void g_free(ograph* g) {
    int i;
    for (i = 0; i < g->nv; i++) { iv_free(&g->nodes[i].out); iv_free(&g->nodes[i].in); }
    free(g->nodes); free(g->edges);
}

/* AlnGraphBoost(const std::string&) :16-39 and AlnGraphBoost(size_t) :41-62 */
static void g_init(ograph* g, uint32_t blen, const char* backbone) {
    uint32_t i;
    memset(g, 0, sizeof *g);
    for (i = 0; i < blen + 2; i++) g_add_vertex(g);
    for (i = 0; i < blen + 1; i++) g_add_edge(g, (int)i, (int)i + 1);     /* count 0 */
    g->enter = 0;
    g->nodes[0].base = '^'; g->nodes[0].backbone = 1;
    for (i = 1; i <= blen; i++) {
        onode* n = &g->nodes[i];
        n->backbone = 1; n->weight = 1;
        n->base = backbone ? backbone[i - 1] : 'N';
        n->bbmap = (int)i;
    }
    g->exit_ = (int)blen + 1;
    g->nodes[g->exit_].base = '$'; g->nodes[g->exit_].backbone = 1;
}

/* addEdge :109-127 */
static void g_aln_edge(ograph* g, int u, int v) {
    int i, exists = 0;
    for (i = 0; i < g->nodes[v].in.n; i++) {
        oedge* e = &g->edges[g->nodes[v].in.v[i]];
        if (e->src == u) { e->count++; exists = 1; }
    }
    if (!exists) { int e = g_add_edge(g, u, v); g->edges[e].count++; }
}

/* addAln :64-107.  Returns 0, or -1 when a match/deletion column indexes a vertex that
 * does not exist (the reference reads out of bounds there: undefined behaviour). */
static int g_add_aln(ograph* g, const char* q, const char* t, int64_t len, uint32_t start) {
    uint32_t bbPos = start;
    int prev = g->enter;
    int64_t i;
    for (i = 0; i < len; i++) {
        char qb = q[i], tb = t[i];
        int cur = (int)bbPos;
        if (qb == tb) {
            if (cur >= g->nv) return -1;
            g->nodes[g->nodes[cur].bbmap].coverage++;
            g->nodes[g->nodes[cur].bbmap].base = tb;
            g->nodes[cur].weight++;
            g_aln_edge(g, prev, cur);
            bbPos++;
            prev = cur;
        } else if (qb == '-' && tb != '-') {
            if (cur >= g->nv) return -1;
            g->nodes[g->nodes[cur].bbmap].coverage++;
            g->nodes[g->nodes[cur].bbmap].base = tb;
            bbPos++;
        } else if (qb != '-' && tb == '-') {
            int nv = g_add_vertex(g);
            g->nodes[nv].base = qb;
            g->nodes[nv].weight++;
            g->nodes[nv].bbmap = (int)bbPos;
            g_aln_edge(g, prev, nv);
            prev = nv;
        }
    }
    g_aln_edge(g, prev, g->exit_);
    return 0;
}

/* markForReaper :269-273 */
static void g_reap(ograph* g, int n) { g->nodes[n].deleted = 1; g_clear_vertex(g, n); }

/* std::map<char, std::vector<VtxDesc>> as a tiny ordered assoc (keys compared as `char`) */
typedef struct { char key[256]; ivec grp[256]; int n; } groups;
static ivec* groups_at(groups* m, char k) {
    int i, j;
    for (i = 0; i < m->n; i++) if (m->key[i] == k) return &m->grp[i];
    for (i = 0; i < m->n; i++) if (k < m->key[i]) break;
    for (j = m->n; j > i; j--) { m->key[j] = m->key[j - 1]; m->grp[j] = m->grp[j - 1]; }
    m->key[i] = k; memset(&m->grp[i], 0, sizeof(ivec)); m->n++;
    return &m->grp[i];
}
static void groups_free(groups* m) { int i; for (i = 0; i < m->n; i++) iv_free(&m->grp[i]); }

/* mergeInNodes :162-215 (recursive) */
static void g_merge_in(ograph* g, int n) {
    groups* m = (groups*)calloc(1, sizeof(groups));
    int i, k;
    for (i = 0; i < g->nodes[n].in.n; i++) {
        int in_node = g->edges[g->nodes[n].in.v[i]].src;
        if (g->nodes[in_node].out.n == 1) iv_push(groups_at(m, g->nodes[in_node].base), in_node);
    }
    for (k = 0; k < m->n; k++) {
        ivec* nodes = &m->grp[k];
        int an, an_out, x;
        if (nodes->n <= 1) continue;
        an = nodes->v[0];
        an_out = g->nodes[an].out.v[0];                 /* *anoi, taken once */
        for (x = 1; x < nodes->n; x++) {
            int ni = nodes->v[x];
            g->edges[an_out].count += g->edges[g->nodes[ni].out.v[0]].count;
            g->nodes[an].weight += g->nodes[ni].weight;
        }
        for (x = 1; x < nodes->n; x++) {
            int ni = nodes->v[x], y;
            for (y = 0; y < g->nodes[ni].in.n; y++) {
                int ie = g->nodes[ni].in.v[y];
                int n1 = g->edges[ie].src;
                int e = g_find_edge(g, n1, an);
                if (e >= 0) {
                    g->edges[e].count += g->edges[ie].count;
                } else {
                    e = g_add_edge(g, n1, an);
                    g->edges[e].count = g->edges[ie].count;
                    g->edges[e].visited = g->edges[ie].visited;
                }
            }
            g_reap(g, ni);
        }
        g_merge_in(g, an);
    }
    groups_free(m); free(m);
}

/* mergeOutNodes :217-267 (not recursive) */
static void g_merge_out(ograph* g, int n) {
    groups* m = (groups*)calloc(1, sizeof(groups));
    int i, k;
    for (i = 0; i < g->nodes[n].out.n; i++) {
        int out_node = g->edges[g->nodes[n].out.v[i]].dst;
        if (g->nodes[out_node].in.n == 1) iv_push(groups_at(m, g->nodes[out_node].base), out_node);
    }
    for (k = 0; k < m->n; k++) {
        ivec* nodes = &m->grp[k];
        int an, an_in, x;
        if (nodes->n <= 1) continue;
        an = nodes->v[0];
        an_in = g->nodes[an].in.v[0];
        for (x = 1; x < nodes->n; x++) {
            int ni = nodes->v[x];
            g->edges[an_in].count += g->edges[g->nodes[ni].in.v[0]].count;
            g->nodes[an].weight += g->nodes[ni].weight;
        }
        for (x = 1; x < nodes->n; x++) {
            int ni = nodes->v[x], y;
            for (y = 0; y < g->nodes[ni].out.n; y++) {
                int oe = g->nodes[ni].out.v[y];
                int n2 = g->edges[oe].dst;
                int e = g_find_edge(g, an, n2);
                if (e >= 0) {
                    g->edges[e].count += g->edges[oe].count;
                } else {
                    e = g_add_edge(g, an, n2);
                    g->edges[e].count = g->edges[oe].count;
                    g->edges[e].visited = g->edges[oe].visited;
                }
            }
            g_reap(g, ni);
        }
    }
    groups_free(m); free(m);
}

/* mergeNodes :129-160 */
static void g_merge_nodes(ograph* g) {
    ivec queue = {0, 0, 0};
    int head = 0;
    iv_push(&queue, g->enter);
    while (head < queue.n) {
        int u = queue.v[head++], i;
        g_merge_in(g, u);
        g_merge_out(g, u);
        for (i = 0; i < g->nodes[u].out.n; i++) {
            int e = g->nodes[u].out.v[i], v, j, not_visited = 0;
            g->edges[e].visited = 1;
            v = g->edges[e].dst;
            for (j = 0; j < g->nodes[v].in.n; j++)
                if (!g->edges[g->nodes[v].in.v[j]].visited) not_visited++;
            if (not_visited == 0) iv_push(&queue, v);
        }
    }
    iv_free(&queue);
}

/* bestPath :375-459.  Fills path_base/path_weight (caller frees) and returns the path length
 * (enter and exit included, as in the reference's std::vector<AlnNode>). */
static int g_best_path(ograph* g, char** path_base, int** path_weight) {
    int i, head = 0, np = 0, prev;
    float* score = (float*)calloc((size_t)g->nv, sizeof(float));     /* nodeScore[], absent = 0.0f */
    int* best = (int*)malloc(sizeof(int) * (size_t)g->nv);           /* bestNodeScoreEdge, -1 = absent */
    ivec queue = {0, 0, 0};
    char* pb = (char*)malloc((size_t)g->nv + 2);
    int* pw = (int*)malloc(sizeof(int) * ((size_t)g->nv + 2));
    for (i = 0; i < g->ne; i++) g->edges[i].visited = 0;
    for (i = 0; i < g->nv; i++) best[i] = -1;

    iv_push(&queue, g->exit_);
    score[g->exit_] = 0.0f;
    while (head < queue.n) {
        int n = queue.v[head++];
        int found = 0, best_e = -1;
        float best_score = -FLT_MAX;
        for (i = 0; i < g->nodes[n].out.n; i++) {
            int e = g->nodes[n].out.v[i];
            int o = g->edges[e].dst;
            onode* on = &g->nodes[o];
            float new_score, sc = score[o];
            if (on->backbone && on->weight == 1) {
                new_score = sc - 10.0f;
            } else {
                onode* bb = &g->nodes[on->bbmap];
                new_score = g->edges[e].count - bb->coverage * 0.5f + sc;
            }
            if (new_score > best_score) { best_score = new_score; best_e = e; found = 1; }
        }
        if (found) { score[n] = best_score; best[n] = best_e; }
        for (i = 0; i < g->nodes[n].in.n; i++) {
            int e = g->nodes[n].in.v[i], s, j, not_visited = 0;
            g->edges[e].visited = 1;
            s = g->edges[e].src;
            for (j = 0; j < g->nodes[s].out.n; j++)
                if (!g->edges[g->nodes[s].out.v[j]].visited) not_visited++;
            if (not_visited == 0) iv_push(&queue, s);
        }
    }
    prev = g->enter;
    for (;;) {
        pb[np] = g->nodes[prev].base; pw[np] = g->nodes[prev].weight; np++;
        if (best[prev] < 0 || np > g->nv) break;      /* np > nv: cycle guard (reference would spin) */
        prev = g->edges[best[prev]].dst;
    }
    free(score); free(best); iv_free(&queue);
    *path_base = pb; *path_weight = pw;
    return np;
}

/* ------------------------------------------------------------------ results + driver */
typedef struct { int range0, range1; int64_t off, len; } ofrag;
typedef struct {
    ofrag* frags; int nfrag, capfrag;
    cbuf seqs;      /* fragment sequences, concatenated */
    cbuf single;    /* consensus(minWeight) result */
    cbuf dump;
    int status;     /* 0 ok, -1 alignment ran off the graph */
} oresult;

static void res_push(oresult* r, int r0, int r1, const char* s, int64_t n) {
    if (r->nfrag == r->capfrag) {
        r->capfrag = r->capfrag ? 2 * r->capfrag : 4;
        r->frags = (ofrag*)realloc(r->frags, sizeof(ofrag) * (size_t)r->capfrag);
    }
    r->frags[r->nfrag].range0 = r0; r->frags[r->nfrag].range1 = r1;
    r->frags[r->nfrag].off = r->seqs.n; r->frags[r->nfrag].len = n;
    r->nfrag++;
    cb_putn(&r->seqs, s, n);
}

/* consensus(std::vector<CnsResult>&, minWeight, minLen) :327-373 */
static void g_consensus_all(ograph* g, oresult* r, int min_weight, int64_t min_len) {
    char* pb; int* pw;
    int np = g_best_path(g, &pb, &pw), i, offs = 0, idx = 0, met = 0;
    char* cns = (char*)malloc((size_t)np + 1);
    char eb = g->nodes[g->enter].base, xb = g->nodes[g->exit_].base;
    for (i = 0; i < np; i++) {
        if (pb[i] == eb || pb[i] == xb) continue;
        cns[idx] = pb[i];
        if (!met && pw[i] >= min_weight) { offs = idx; met = 1; }
        else if (met && pw[i] < min_weight) {
            met = 0;
            if ((int64_t)(idx - offs) >= min_len) res_push(r, offs, idx, cns + offs, idx - offs);
        }
        idx++;
    }
    if (met && (int64_t)(idx - offs) >= min_len) res_push(r, offs, idx, cns + offs, idx - offs);
    free(cns); free(pb); free(pw);
}

/* consensus(minWeight) :285-325 -- the single longest qualifying run */
static void g_consensus_single(ograph* g, oresult* r, int min_weight) {
    char* pb; int* pw;
    int np = g_best_path(g, &pb, &pw), i, offs = 0, best_offs = 0, length = 0, idx = 0, met = 0;
    char* cns = (char*)malloc((size_t)np + 1);
    char eb = g->nodes[g->enter].base, xb = g->nodes[g->exit_].base;
    for (i = 0; i < np; i++) {
        if (pb[i] == eb || pb[i] == xb) continue;
        cns[idx] = pb[i];
        if (!met && pw[i] >= min_weight) { offs = idx; met = 1; }
        else if (met && pw[i] < min_weight) {
            if ((idx - offs) > length) { best_offs = offs; length = idx - offs; }
            met = 0;
        }
        idx++;
    }
    if (met && (idx - offs) > length) { best_offs = offs; length = idx - offs; }
    cb_putn(&r->single, cns + best_offs, length);
    free(cns); free(pb); free(pw);
}

static void g_dump(ograph* g, cbuf* out) {
    char buf[128];
    int v, i;
    for (v = 0; v < g->nv; v++) {
        onode* n = &g->nodes[v];
        snprintf(buf, sizeof buf, "%d %c %d %d %d %d |", v, n->base, n->weight, n->coverage, (int)n->backbone, (int)n->deleted);
        cb_putn(out, buf, (int64_t)strlen(buf));
        for (i = 0; i < n->out.n; i++) {
            oedge* e = &g->edges[n->out.v[i]];
            snprintf(buf, sizeof buf, " %d:%d:%d", e->dst, e->count, (int)e->visited);
            cb_putn(out, buf, (int64_t)strlen(buf));
        }
        cb_putn(out, " |", 2);
        for (i = 0; i < n->in.n; i++) {
            oedge* e = &g->edges[n->in.v[i]];
            snprintf(buf, sizeof buf, " %d:%d:%d", e->src, e->count, (int)e->visited);
            cb_putn(out, buf, (int64_t)strlen(buf));
        }
        cb_putc(out, '\n');
    }
}

/* The Consensus functor for one target: main.cpp:130-138 / dazcon.cpp:76-89.
 * mode bits: 1 normalizeGaps, 2 trimAln, 4 mergeNodes, 8 single-string consensus,
 *            16 dump graph before merge, 32 dump after merge (same text as oracle/ref_capi.cpp).
 * backbone == NULL -> size_t ctor. */
void* orc_consensus(uint32_t tlen, const char* backbone, int n_aln, const char* const* q,
                    const char* const* t, const int64_t* len, const uint32_t* start, int mode,
                    int trim, int min_weight, int64_t min_len) {
    oresult* r = (oresult*)calloc(1, sizeof(oresult));
    ograph g;
    int i;
    g_init(&g, tlen, backbone);
    for (i = 0; i < n_aln && r->status == 0; i++) {
        const char *qq = q[i], *tt = t[i];
        char *qn = 0, *tn = 0;
        int64_t n = len[i], loffs = 0;
        uint32_t st = start[i];
        if (n < min_len) continue;                       /* raw qstr.length() filter */
        if (mode & 1) {
            qn = (char*)malloc((size_t)(2 * n + 1)); tn = (char*)malloc((size_t)(2 * n + 1));
            n = orc_normalize_gaps(qq, tt, n, qn, tn);
            qq = qn; tt = tn;
        }
        if (mode & 2) n = orc_trim(tt, n, trim, &st, &loffs);
        if (g_add_aln(&g, qq + loffs, tt + loffs, n, st) != 0) r->status = -1;
        free(qn); free(tn);
    }
    if (r->status == 0) {
        if (mode & 16) g_dump(&g, &r->dump);
        if (mode & 4) g_merge_nodes(&g);
        if (mode & 32) g_dump(&g, &r->dump);
        if (mode & 8) g_consensus_single(&g, r, min_weight);
        else g_consensus_all(&g, r, min_weight, min_len);
    }
    g_free(&g);
    return r;
}

int orc_result_status(void* h) { return ((oresult*)h)->status; }
int orc_result_count(void* h) { return ((oresult*)h)->nfrag; }
int orc_result_range(void* h, int i, int which) { oresult* r = (oresult*)h; return which ? r->frags[i].range1 : r->frags[i].range0; }
int64_t orc_result_seq(void* h, int i, const char** p) { oresult* r = (oresult*)h; *p = r->seqs.p + r->frags[i].off; return r->frags[i].len; }
int64_t orc_result_single(void* h, const char** p) { oresult* r = (oresult*)h; *p = r->single.p ? r->single.p : ""; return r->single.n; }
int64_t orc_result_dump(void* h, const char** p) { oresult* r = (oresult*)h; *p = r->dump.p ? r->dump.p : ""; return r->dump.n; }
void orc_result_free(void* h) { oresult* r = (oresult*)h; free(r->frags); free(r->seqs.p); free(r->single.p); free(r->dump.p); free(r); }
