/* TEST INFRASTRUCTURE — CPU oracle, plain C restatement of the reference's hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this. */
#ifndef BRWT_ORACLE_H_
#define BRWT_ORACLE_H_
#include <stddef.h>
#include <stdint.h>

typedef struct orc_bits orc_bits;     /* plain bit vector + prefix popcounts */
typedef struct orc_brwt orc_brwt;     /* BRWT node tree */
typedef struct orc_colmajor orc_colmajor;

/* bit vectors: kind 0 = bit_vector_stat, 1 = bit_vector_sd, 2 = bit_vector_rrr<63> */
orc_bits *orc_bv_load(int kind, const uint8_t *buf, size_t len, size_t *consumed);
orc_bits *orc_bv_from_words(const uint64_t *words, uint64_t n_bits);
void orc_bv_free(orc_bits *v);
uint64_t orc_bv_size(const orc_bits *v);
uint64_t orc_bv_num_set_bits(const orc_bits *v);
int orc_bv_inverted(const orc_bits *v);
uint64_t orc_bv_rank1(const orc_bits *v, uint64_t id);
uint64_t orc_bv_select1(const orc_bits *v, uint64_t i);
int orc_bv_get(const orc_bits *v, uint64_t id);

orc_brwt *orc_brwt_load(const uint8_t *buf, size_t len, size_t *consumed);
void orc_brwt_free(orc_brwt *m);
uint64_t orc_brwt_num_rows(const orc_brwt *m);
uint64_t orc_brwt_num_columns(const orc_brwt *m);
uint64_t orc_brwt_num_relations(const orc_brwt *m);
uint64_t orc_brwt_num_nodes(const orc_brwt *m);
uint64_t orc_brwt_total_column_size(const orc_brwt *m);
int orc_brwt_get(const orc_brwt *m, uint64_t row, uint64_t col);
uint64_t orc_brwt_get_row(const orc_brwt *m, uint64_t row, uint64_t *out, uint64_t cap);
uint64_t orc_brwt_get_column(const orc_brwt *m, uint64_t col, uint64_t *out, uint64_t cap);
/* number of (node,row) touches of get_row(row): V_t of SURVEY.md 8(d) */
uint64_t orc_brwt_row_touches(const orc_brwt *m, uint64_t row);
/* timing helper: get_row over rows[0..n), returns a checksum */
uint64_t orc_brwt_scan_rows(const orc_brwt *m, const uint64_t *rows, uint64_t n);

orc_colmajor *orc_colmajor_load(const uint8_t *buf, size_t len, size_t *consumed);
void orc_colmajor_free(orc_colmajor *m);
uint64_t orc_colmajor_num_rows(const orc_colmajor *m);
uint64_t orc_colmajor_num_columns(const orc_colmajor *m);
uint64_t orc_colmajor_num_relations(const orc_colmajor *m);
int orc_colmajor_get(const orc_colmajor *m, uint64_t row, uint64_t col);
uint64_t orc_colmajor_get_row(const orc_colmajor *m, uint64_t row, uint64_t *out, uint64_t cap);
uint64_t orc_colmajor_get_column(const orc_colmajor *m, uint64_t col, uint64_t *out, uint64_t cap);

/* portable synthetic matrix of SURVEY.md 8(d) */
uint64_t orc_mix64(uint64_t x);
uint64_t orc_synth_row(uint64_t n_cols, uint64_t threshold, uint64_t seed, uint64_t row, uint64_t *out);
void orc_synth_bits(uint64_t n_bits, uint64_t threshold, uint64_t seed, uint64_t *words);

#endif
