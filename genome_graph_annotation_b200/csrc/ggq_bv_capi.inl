// placeholder C ABI for bit vectors
extern "C" {
#define NI return ggq::fail(GGQ_ERR_INTERNAL, "bit vectors: not implemented yet")
int ggq_bv_load_buffer(int, const uint8_t *, size_t, size_t *, int, ggq_bv **) { NI; }
int ggq_bv_build(int, const uint64_t *, uint64_t, int, ggq_bv **) { NI; }
int ggq_bv_synthesize(int, uint64_t, uint64_t, uint64_t, int, ggq_bv **) { NI; }
void ggq_bv_free(ggq_bv *) {}
uint64_t ggq_bv_size(const ggq_bv *) { return 0; }
uint64_t ggq_bv_num_set_bits(const ggq_bv *) { return 0; }
int ggq_bv_is_inverted(const ggq_bv *) { return 0; }
uint64_t ggq_bv_device_bytes(const ggq_bv *) { return 0; }
int ggq_bv_rank1(const ggq_bv *, const uint64_t *, uint64_t, uint64_t *) { NI; }
int ggq_bv_select1(const ggq_bv *, const uint64_t *, uint64_t, uint64_t *) { NI; }
int ggq_bv_get(const ggq_bv *, const uint64_t *, uint64_t, uint8_t *) { NI; }
int ggq_bv_rank1_device(const ggq_bv *, const uint64_t *, uint64_t, uint64_t *, void *) { NI; }
int ggq_bv_select1_device(const ggq_bv *, const uint64_t *, uint64_t, uint64_t *, void *) { NI; }
int ggq_bv_get_device(const ggq_bv *, const uint64_t *, uint64_t, uint8_t *, void *) { NI; }
#undef NI
}
