/* TEST INFRASTRUCTURE — CPU oracle.  NOT part of the product; nothing under
 * genome_graph_annotation_b200/ may link, load or call this file.
 *
 * Plain-C restatement of the reference's query hot path, written for obviousness, not speed:
 * vectors are decoded to plain bits, rank is a prefix-popcount table, select a binary search on it.
 * Every function cites the reference code it follows (paths relative to /root/reference).
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file against the golden vectors generated
 * from the reference's own code (tests/golden/, made by tests/golden/make_golden.py through
 * oracle/_ref/libggref.so, i.e. the unmodified reference compiled by oracle/Makefile), and, when
 * oracle/_ref is present, directly against that library on randomized inputs.  The reference's own
 * unit tests for the path (94 gtest cases) pass against oracle/_ref (tests/test_oracle_ref.py).
 */
#include "brwt_oracle.h"

#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------ */
/* byte reader                                                                                */
/* ------------------------------------------------------------------------------------------ */
typedef struct { const uint8_t *p; size_t len, pos; int bad; } rd_t;

static void rd_bytes(rd_t *r, void *dst, size_t n) {
    if (r->bad || n > r->len - r->pos) { r->bad = 1; memset(dst, 0, n); return; }
    memcpy(dst, r->p + r->pos, n);
    r->pos += n;
}
/* libmaus2 NumberSerialisation: 8 bytes, most significant first (common/serialization.cpp:27-36) */
static uint64_t rd_be64(rd_t *r) {
    uint8_t b[8]; uint64_t v = 0; int i;
    rd_bytes(r, b, 8);
    for (i = 0; i < 8; ++i) v = (v << 8) | b[i];
    return v;
}
static uint64_t rd_le64(rd_t *r) { uint64_t v = 0; rd_bytes(r, &v, 8); return v; }
static uint8_t rd_u8(rd_t *r) { uint8_t v = 0; rd_bytes(r, &v, 1); return v; }

/* sdsl::int_vector<t_width>::load (sdsl/int_vector.hpp:1582-1599): LE u64 size in bits,
 * a width byte only for int_vector<0>, then ceil(bits/64) LE words. */
typedef struct { uint64_t bits; uint8_t width; uint64_t *w; uint64_t n_words; } iv_t;

static void iv_free(iv_t *v) { free(v->w); v->w = NULL; }
static int iv_load(rd_t *r, iv_t *v, int has_width, uint8_t fixed) {
    v->bits = rd_le64(r);
    v->width = has_width ? rd_u8(r) : fixed;
    v->n_words = (v->bits + 63) / 64;
    v->w = NULL;
    if (r->bad || v->n_words > (r->len - r->pos) / 8) { r->bad = 1; return 0; }
    v->w = (uint64_t *)calloc(v->n_words + 2, 8);
    rd_bytes(r, v->w, v->n_words * 8);
    return !r->bad;
}
static uint64_t iv_bits(const iv_t *v, uint64_t pos, unsigned len) {
    uint64_t x; unsigned off = (unsigned)(pos & 63);
    if (!len) return 0;
    x = v->w[pos >> 6] >> off;
    if (off + len > 64) x |= v->w[(pos >> 6) + 1] << (64 - off);
    if (len < 64) x &= ((uint64_t)1 << len) - 1;
    return x;
}
static uint64_t iv_get(const iv_t *v, uint64_t i) { return iv_bits(v, i * v->width, v->width); }
static uint64_t iv_size(const iv_t *v) { return v->width ? v->bits / v->width : 0; }

/* ------------------------------------------------------------------------------------------ */
/* plain bit vector with rank / select                                                        */
/* ------------------------------------------------------------------------------------------ */
struct orc_bits {
    uint64_t n;        /* size in bits */
    uint64_t *w;       /* ceil(n/64)+1 words */
    uint64_t *pre;     /* pre[k] = ones in words [0, k) */
    uint64_t ones;
    int inverted;      /* bit_vector_sd::is_inverted() of the stream this came from */
};

static orc_bits *bits_new(uint64_t n) {
    orc_bits *v = (orc_bits *)calloc(1, sizeof(orc_bits));
    v->n = n;
    v->w = (uint64_t *)calloc((n + 63) / 64 + 2, 8);
    return v;
}
static void bits_set(orc_bits *v, uint64_t i) { v->w[i >> 6] |= (uint64_t)1 << (i & 63); }
static void bits_finish(orc_bits *v) {
    uint64_t k, nw = (v->n + 63) / 64;
    if (v->n & 63) v->w[nw - 1] &= ((uint64_t)1 << (v->n & 63)) - 1;
    v->pre = (uint64_t *)calloc(nw + 2, 8);
    for (k = 0; k < nw; ++k) v->pre[k + 1] = v->pre[k] + (uint64_t)__builtin_popcountll(v->w[k]);
    v->ones = v->pre[nw];
}
void orc_bv_free(orc_bits *v) { if (v) { free(v->w); free(v->pre); free(v); } }
uint64_t orc_bv_size(const orc_bits *v) { return v->n; }
uint64_t orc_bv_num_set_bits(const orc_bits *v) { return v->ones; }
int orc_bv_inverted(const orc_bits *v) { return v->inverted; }

/* operator[]: common/bit_vector.cpp:296-299 (stat), :708-711 (sd: stored bit != inverted), :878-882 (rrr) */
int orc_bv_get(const orc_bits *v, uint64_t id) { return (int)((v->w[id >> 6] >> (id & 63)) & 1); }

/* rank1(id) = number of ones in [0, id], with id >= size clamped to the whole vector:
 * common/bit_vector.cpp:281-286 (stat), :690-695 (sd), :856-861 (rrr) -> sdsl rank(min(id+1, size)). */
uint64_t orc_bv_rank1(const orc_bits *v, uint64_t id) {
    uint64_t idx = id >= v->n ? v->n : id + 1;   /* exclusive end */
    uint64_t r = v->pre[idx >> 6];
    if (idx & 63) r += (uint64_t)__builtin_popcountll(v->w[idx >> 6] & (((uint64_t)1 << (idx & 63)) - 1));
    return r;
}

/* select1(i) = position of the i-th one, i in [1, ones]:
 * common/bit_vector.cpp:288-295 (stat), :697-702 (sd; select0 of the stored vector when inverted), :863-869 (rrr) */
uint64_t orc_bv_select1(const orc_bits *v, uint64_t i) {
    uint64_t lo = 0, hi = (v->n + 63) / 64, x;
    while (hi - lo > 1) {               /* last word k with pre[k] < i */
        uint64_t mid = (lo + hi) / 2;
        if (v->pre[mid] < i) lo = mid; else hi = mid;
    }
    x = v->w[lo];
    i -= v->pre[lo];
    while (--i) x &= x - 1;
    return lo * 64 + (uint64_t)__builtin_ctzll(x);
}

orc_bits *orc_bv_from_words(const uint64_t *words, uint64_t n_bits) {
    orc_bits *v = bits_new(n_bits);
    memcpy(v->w, words, ((n_bits + 63) / 64) * 8);
    bits_finish(v);
    return v;
}

/* ---- rrr_vector<63> ------------------------------------------------------------------------ */
static uint64_t binom[64][64];
static uint8_t rrr_space[64];
static int binom_ready = 0;
static void binom_init(void) {
    int n, k;
    if (binom_ready) return;
    /* sdsl/rrr_helper.hpp:186-199: Pascal triangle; :267-272: space[k] = 0 if C(63,k)==1 else hi(C)+1 */
    for (n = 0; n < 64; ++n) for (k = 0; k < 64; ++k) binom[n][k] = 0;
    for (n = 0; n < 64; ++n) binom[n][0] = 1;
    for (n = 1; n < 64; ++n) for (k = 1; k <= n; ++k) binom[n][k] = binom[n - 1][k - 1] + binom[n - 1][k];
    for (k = 0; k < 64; ++k)
        rrr_space[k] = binom[63][k] == 1 ? 0 : (uint8_t)(64 - __builtin_clzll(binom[63][k]));
    binom_ready = 1;
}
/* inverse of rrr_helper::bin_to_nr (sdsl/rrr_helper.hpp:295-312), same walk as decode_int (:374-405) */
static uint64_t rrr_block(unsigned k, uint64_t nr) {
    uint64_t res = 0; unsigned nn = 63, i = 0;
    if (k == 63) return ((uint64_t)1 << 63) - 1;
    if (k == 0) return 0;
    while (k > 1) {
        if (nr >= binom[nn - 1][k]) { nr -= binom[nn - 1][k]; --k; res |= (uint64_t)1 << i; }
        --nn; ++i;
    }
    res |= (uint64_t)1 << (63 - nr - 1);
    return res;
}
/* rrr_vector::load (sdsl/rrr_vector.hpp:365-373) + block walk of operator[] (:257-277) */
static orc_bits *rrr_load(rd_t *r) {
    iv_t bt, btnr, btnrp, rank, inv;
    uint64_t m_size, n_blocks, b, pos = 0;
    orc_bits *v = NULL;
    binom_init();
    m_size = rd_le64(r);
    iv_load(r, &bt, 1, 0); iv_load(r, &btnr, 0, 1); iv_load(r, &btnrp, 1, 0);
    iv_load(r, &rank, 1, 0); iv_load(r, &inv, 0, 1);
    if (!r->bad) {
        v = bits_new(m_size);
        n_blocks = iv_size(&bt);
        if (bt.bits == 0) n_blocks = 0;
        for (b = 0; b < n_blocks && !r->bad; ++b) {
            unsigned stored, k, sp; uint64_t start = b * 63, blockbits, j;
            if (b % 32 == 0) pos = iv_get(&btnrp, b / 32);
            stored = (unsigned)iv_get(&bt, b);
            if (stored > 63) { r->bad = 1; break; }
            k = ((inv.w[(b / 32) >> 6] >> ((b / 32) & 63)) & 1) ? 63 - stored : stored;  /* :261-263 */
            sp = rrr_space[stored];
            if (pos + sp > btnr.bits) { r->bad = 1; break; }
            blockbits = sp ? rrr_block(k, iv_bits(&btnr, pos, sp)) : (k == 63 ? ((uint64_t)1 << 63) - 1 : 0);
            pos += sp;
            for (j = 0; j < 63 && start + j < m_size; ++j)
                if ((blockbits >> j) & 1) bits_set(v, start + j);
        }
        if (r->bad) { orc_bv_free(v); v = NULL; } else bits_finish(v);
    }
    iv_free(&bt); iv_free(&btnr); iv_free(&btnrp); iv_free(&rank); iv_free(&inv);
    return v;
}

/* ---- select_support_mcl stream (sdsl/select_support_mcl.hpp:425-494): parsed only to be skipped */
static void mcl_skip(rd_t *r) {
    uint64_t arg_cnt = rd_le64(r), sb, i; iv_t t;
    if (r->bad || !arg_cnt) return;
    sb = (arg_cnt + 4095) >> 12;
    if (iv_load(r, &t, 1, 0)) iv_free(&t); else return;
    if (iv_load(r, &t, 0, 1)) iv_free(&t); else return;
    for (i = 0; i < sb && !r->bad; ++i) { if (iv_load(r, &t, 1, 0)) iv_free(&t); }
}

/* ---- bit_vector_sd: sd_vector<> (sdsl/sd_vector.hpp:404-432) + inverted byte (common/bit_vector.cpp:725-745).
 * i-th one of the stored vector = ((select1_high(i+1) - i) << wl) | low[i]   (sd_vector.hpp:544-553);
 * logical bit = stored bit != inverted (common/bit_vector.cpp:708-711). */
static orc_bits *sd_load(rd_t *r) {
    uint64_t size = rd_le64(r), m, i = 0, hp;
    uint8_t wl = rd_u8(r), invb;
    iv_t low, high;
    orc_bits *v = NULL;
    iv_load(r, &low, 1, 0); iv_load(r, &high, 0, 1);
    mcl_skip(r); mcl_skip(r);
    invb = rd_u8(r);
    if (!r->bad && invb <= 1 && wl < 64) {
        uint64_t k, nw;
        v = bits_new(size);
        m = iv_size(&low);
        if (low.bits == 0) m = 0;
        for (hp = 0; hp < high.bits && i < m; ++hp) {
            if ((high.w[hp >> 6] >> (hp & 63)) & 1) {
                uint64_t p = ((hp - i) << wl) | (wl ? iv_get(&low, i) : 0);
                if (p >= size) { r->bad = 1; break; }
                bits_set(v, p);
                ++i;
            }
        }
        if (i != m) r->bad = 1;
        if (invb) {
            nw = (size + 63) / 64;
            for (k = 0; k < nw; ++k) v->w[k] = ~v->w[k];
        }
        v->inverted = invb;
        if (r->bad) { orc_bv_free(v); v = NULL; } else bits_finish(v);
    } else {
        r->bad = 1;
    }
    iv_free(&low); iv_free(&high);
    return v;
}

/* ---- bit_vector_stat (common/bit_vector.cpp:346-385): bit_vector, BE popcount, rank_support_v5, select_support_mcl;
 * missing supports are tolerated like the reference does (:367-380). */
static orc_bits *stat_load(rd_t *r) {
    iv_t bv, rk; orc_bits *v;
    if (!iv_load(r, &bv, 0, 1)) return NULL;
    v = bits_new(bv.bits);
    memcpy(v->w, bv.w, bv.n_words * 8);
    iv_free(&bv);
    bits_finish(v);
    {
        rd_t save = *r;
        (void)rd_be64(r);
        if (iv_load(r, &rk, 0, 64)) iv_free(&rk);
        mcl_skip(r);
        if (r->bad) { *r = save; r->pos = r->len; r->bad = 0; }
    }
    return v;
}

orc_bits *orc_bv_load(int kind, const uint8_t *buf, size_t len, size_t *consumed) {
    rd_t r = { buf, len, 0, 0 };
    orc_bits *v = kind == 0 ? stat_load(&r) : (kind == 1 ? sd_load(&r) : rrr_load(&r));
    if (consumed) *consumed = r.pos;
    return v;
}

/* ------------------------------------------------------------------------------------------ */
/* BRWT                                                                                       */
/* ------------------------------------------------------------------------------------------ */
typedef struct node {
    /* utils::RangePartition (common/utils.hpp:440-479) */
    uint32_t n_groups;
    uint32_t **partition;   /* partition_[g][r] */
    uint32_t *group_size;
    uint32_t *groups, *ranks;
    uint64_t n_cols;
    orc_bits *nonzero_rows;
    uint32_t n_children;
    struct node **children;
} node_t;

struct orc_brwt { node_t *root; };

static void node_free(node_t *n) {
    uint32_t i;
    if (!n) return;
    for (i = 0; i < n->n_children; ++i) node_free(n->children[i]);
    if (n->partition) for (i = 0; i < n->n_groups; ++i) free(n->partition[i]);
    free(n->partition); free(n->group_size); free(n->groups); free(n->ranks); free(n->children);
    orc_bv_free(n->nonzero_rows);
    free(n);
}

/* BRWT::load (annotation/hierarchical_annotation/BRWT.cpp:87-111), RangePartition::load (common/utils.cpp:702-715) */
static node_t *node_load(rd_t *r, int depth) {
    node_t *n = (node_t *)calloc(1, sizeof(node_t));
    uint64_t g, total = 0, nc;
    if (depth > 4096) r->bad = 1;
    n->n_groups = (uint32_t)rd_be64(r);
    if (r->bad || n->n_groups > r->len) { r->bad = 1; n->n_groups = 0; node_free(n); return NULL; }
    n->partition = (uint32_t **)calloc(n->n_groups + 1, sizeof(uint32_t *));
    n->group_size = (uint32_t *)calloc(n->n_groups + 1, 4);
    for (g = 0; g < n->n_groups && !r->bad; ++g) {
        iv_t v; uint64_t k, i;
        if (!iv_load(r, &v, 1, 0)) break;
        k = iv_size(&v);
        if (v.bits == 0) k = 0;
        if (!k) r->bad = 1;                       /* groups must not be empty: utils.cpp:655-659 */
        n->partition[g] = (uint32_t *)calloc(k + 1, 4);
        n->group_size[g] = (uint32_t)k;
        for (i = 0; i < k; ++i) n->partition[g][i] = (uint32_t)iv_get(&v, i);
        total += k;
        iv_free(&v);
    }
    if (!r->bad) {                                 /* initialize_groups_and_ranks: utils.cpp:651-680 */
        uint64_t i;
        n->n_cols = total;
        n->groups = (uint32_t *)malloc((total + 1) * 4);
        n->ranks = (uint32_t *)malloc((total + 1) * 4);
        memset(n->groups, 0xff, (total + 1) * 4);
        for (g = 0; g < n->n_groups; ++g)
            for (i = 0; i < n->group_size[g]; ++i) {
                uint32_t c = n->partition[g][i];
                if (c >= total || n->groups[c] != 0xffffffffu) { r->bad = 1; break; }
                n->groups[c] = (uint32_t)g; n->ranks[c] = (uint32_t)i;
            }
    }
    if (!r->bad) n->nonzero_rows = rrr_load(r);
    if (!r->bad) {
        nc = rd_be64(r);
        if (r->bad || (nc && nc != n->n_groups)) r->bad = 1;    /* BRWT.cpp:106-107 */
        else {
            n->children = (node_t **)calloc(nc + 1, sizeof(node_t *));
            for (g = 0; g < nc && !r->bad; ++g) {
                n->children[g] = node_load(r, depth + 1);
                if (n->children[g]) n->n_children = (uint32_t)(g + 1);
            }
        }
    }
    if (r->bad) { node_free(n); return NULL; }
    return n;
}

orc_brwt *orc_brwt_load(const uint8_t *buf, size_t len, size_t *consumed) {
    rd_t r = { buf, len, 0, 0 };
    node_t *root = node_load(&r, 0);
    orc_brwt *m;
    if (consumed) *consumed = r.pos;
    if (!root) return NULL;
    m = (orc_brwt *)calloc(1, sizeof(orc_brwt));
    m->root = root;
    return m;
}
void orc_brwt_free(orc_brwt *m) { if (m) { node_free(m->root); free(m); } }
uint64_t orc_brwt_num_rows(const orc_brwt *m) { return m->root->nonzero_rows->n; }      /* BRWT.hpp:34 */
uint64_t orc_brwt_num_columns(const orc_brwt *m) { return m->root->n_cols; }             /* BRWT.hpp:33 */

/* BRWT::num_relations (BRWT.cpp:130-140) */
static uint64_t node_relations(const node_t *n) {
    uint64_t s = 0; uint32_t i;
    if (!n->n_children) return n->nonzero_rows->ones;
    for (i = 0; i < n->n_children; ++i) s += node_relations(n->children[i]);
    return s;
}
uint64_t orc_brwt_num_relations(const orc_brwt *m) { return node_relations(m->root); }
static uint64_t node_count(const node_t *n) {
    uint64_t s = 1; uint32_t i;
    for (i = 0; i < n->n_children; ++i) s += node_count(n->children[i]);
    return s;
}
uint64_t orc_brwt_num_nodes(const orc_brwt *m) { return node_count(m->root); }          /* BRWT.cpp:160-166 */
static uint64_t node_colsize(const node_t *n) {
    uint64_t s = n->nonzero_rows->n; uint32_t i;
    for (i = 0; i < n->n_children; ++i) s += node_colsize(n->children[i]);
    return s;
}
uint64_t orc_brwt_total_column_size(const orc_brwt *m) { return node_colsize(m->root); }  /* BRWT.cpp:183-191 */

/* BRWT::get (BRWT.cpp:9-24) */
static int node_get(const node_t *n, uint64_t row, uint64_t col) {
    if (!orc_bv_get(n->nonzero_rows, row)) return 0;
    if (!n->n_children) return 1;
    return node_get(n->children[n->groups[col]], orc_bv_rank1(n->nonzero_rows, row) - 1, n->ranks[col]);
}
int orc_brwt_get(const orc_brwt *m, uint64_t row, uint64_t col) { return node_get(m->root, row, col); }

/* BRWT::get_row (BRWT.cpp:26-53): children in stored order, child-local column c -> partition_[i][c].
 * Returns the count; writes into out (capacity cap) in the reference's order. */
static uint64_t node_get_row(const node_t *n, uint64_t row, uint64_t *out, uint64_t cap, uint64_t *touches) {
    uint64_t cnt = 0, idx; uint32_t i;
    if (touches) ++*touches;
    if (!orc_bv_get(n->nonzero_rows, row)) return 0;
    if (!n->n_children) {            /* leaf: utils::arange(0, assignments_.size()) */
        for (idx = 0; idx < n->n_cols; ++idx) { if (cnt < cap) out[cnt] = idx; ++cnt; }
        return cnt;
    }
    idx = orc_bv_rank1(n->nonzero_rows, row) - 1;
    for (i = 0; i < n->n_children; ++i) {
        uint64_t k = node_get_row(n->children[i], idx, out + cnt, cap > cnt ? cap - cnt : 0, touches), j;
        for (j = 0; j < k && cnt + j < cap; ++j) out[cnt + j] = n->partition[i][out[cnt + j]];
        cnt += k;
    }
    return cnt;
}
uint64_t orc_brwt_get_row(const orc_brwt *m, uint64_t row, uint64_t *out, uint64_t cap) {
    return node_get_row(m->root, row, out, cap, NULL);
}
uint64_t orc_brwt_row_touches(const orc_brwt *m, uint64_t row) {
    uint64_t t = 0, buf_cap = m->root->n_cols + 1;
    uint64_t *buf = (uint64_t *)malloc(buf_cap * 8);
    node_get_row(m->root, row, buf, buf_cap, &t);
    free(buf);
    return t;
}
uint64_t orc_brwt_scan_rows(const orc_brwt *m, const uint64_t *rows, uint64_t n) {
    uint64_t cap = m->root->n_cols + 1, sum = 0, t, j;
    uint64_t *buf = (uint64_t *)malloc(cap * 8);
    for (t = 0; t < n; ++t) {
        uint64_t k = node_get_row(m->root, rows[t], buf, cap, NULL);
        sum += k;
        for (j = 0; j < k; ++j) sum += buf[j];
    }
    free(buf);
    return sum;
}

/* BRWT::get_column (BRWT.cpp:55-85) */
static uint64_t node_get_column(const node_t *n, uint64_t col, uint64_t *out, uint64_t cap) {
    const orc_bits *v = n->nonzero_rows;
    uint64_t cnt, i;
    if (!v->ones) return 0;                              /* :58-62 */
    if (!n->n_children) {                                /* :65-71 call_ones */
        cnt = 0;
        for (i = 0; i < v->n; ++i) if (orc_bv_get(v, i)) { if (cnt < cap) out[cnt] = i; ++cnt; }
        return cnt;
    }
    cnt = node_get_column(n->children[n->groups[col]], n->ranks[col], out, cap);   /* :73-74 */
    if (v->ones == v->n) return cnt;                     /* :77-78 */
    for (i = 0; i < cnt && i < cap; ++i) out[i] = orc_bv_select1(v, out[i] + 1);    /* :81-83 */
    return cnt;
}
uint64_t orc_brwt_get_column(const orc_brwt *m, uint64_t col, uint64_t *out, uint64_t cap) {
    return node_get_column(m->root, col, out, cap);
}

/* ------------------------------------------------------------------------------------------ */
/* ColMajorCompressed (annotation/column_compressed/column_major.cpp:27-88)                   */
/* ------------------------------------------------------------------------------------------ */
struct orc_colmajor { uint64_t m; orc_bits **cols; };

orc_colmajor *orc_colmajor_load(const uint8_t *buf, size_t len, size_t *consumed) {
    rd_t r = { buf, len, 0, 0 };
    orc_colmajor *c = (orc_colmajor *)calloc(1, sizeof(orc_colmajor));
    uint64_t j;
    c->m = rd_be64(&r);                                  /* :66 */
    if (r.bad || c->m > len) { free(c); if (consumed) *consumed = r.pos; return NULL; }
    c->cols = (orc_bits **)calloc(c->m + 1, sizeof(orc_bits *));
    for (j = 0; j < c->m; ++j) {
        c->cols[j] = sd_load(&r);
        if (!c->cols[j]) { orc_colmajor_free(c); if (consumed) *consumed = r.pos; return NULL; }
    }
    if (consumed) *consumed = r.pos;
    return c;
}
void orc_colmajor_free(orc_colmajor *c) {
    uint64_t j;
    if (!c) return;
    for (j = 0; j < c->m; ++j) orc_bv_free(c->cols[j]);
    free(c->cols); free(c);
}
uint64_t orc_colmajor_num_columns(const orc_colmajor *c) { return c->m; }
uint64_t orc_colmajor_num_rows(const orc_colmajor *c) { return c->m ? c->cols[0]->n : 0; }   /* :18-25 */
uint64_t orc_colmajor_num_relations(const orc_colmajor *c) {
    uint64_t s = 0, j;
    for (j = 0; j < c->m; ++j) s += c->cols[j]->ones;
    return s;
}
int orc_colmajor_get(const orc_colmajor *c, uint64_t row, uint64_t col) { return orc_bv_get(c->cols[col], row); }
uint64_t orc_colmajor_get_row(const orc_colmajor *c, uint64_t row, uint64_t *out, uint64_t cap) {
    uint64_t cnt = 0, j;                                /* :34-47 ascending column order */
    for (j = 0; j < c->m; ++j) if (orc_bv_get(c->cols[j], row)) { if (cnt < cap) out[cnt] = j; ++cnt; }
    return cnt;
}
uint64_t orc_colmajor_get_column(const orc_colmajor *c, uint64_t col, uint64_t *out, uint64_t cap) {
    const orc_bits *v = c->cols[col];                   /* :49-56 call_ones */
    uint64_t cnt = 0, i;
    for (i = 0; i < v->n; ++i) if (orc_bv_get(v, i)) { if (cnt < cap) out[cnt] = i; ++cnt; }
    return cnt;
}

/* ------------------------------------------------------------------------------------------ */
/* portable synthetic data: splitmix64 finaliser                             */
/* ------------------------------------------------------------------------------------------ */
uint64_t orc_mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL;
    x ^= x >> 27; x *= 0x94d049bb133111ebULL;
    x ^= x >> 31;
    return x;
}
uint64_t orc_synth_row(uint64_t n_cols, uint64_t threshold, uint64_t seed, uint64_t row, uint64_t *out) {
    uint64_t base = seed * 0x9E3779B97F4A7C15ULL + row * n_cols, j, cnt = 0;
    for (j = 0; j < n_cols; ++j) if (orc_mix64(base + j) < threshold) out[cnt++] = j;
    return cnt;
}
void orc_synth_bits(uint64_t n_bits, uint64_t threshold, uint64_t seed, uint64_t *words) {
    uint64_t base = seed * 0x9E3779B97F4A7C15ULL, i;
    memset(words, 0, ((n_bits + 63) / 64) * 8);
    for (i = 0; i < n_bits; ++i) if (orc_mix64(base + i) < threshold) words[i >> 6] |= (uint64_t)1 << (i & 63);
}
