#pragma once
#include "ggq_engine.hpp"
namespace ggq {
struct ColMajorImage {};
inline Matrix *colmajor_load(Reader &, int) { throw std::runtime_error("column-major: not implemented yet"); }
inline void colmajor_rows_device(const Matrix *, Workspace *, const u64 *, u64, int, u64 *) { throw std::runtime_error("column-major: not implemented yet"); }
inline void colmajor_get_launch(const Matrix *, const u64 *, const u64 *, u64, u64, unsigned char *, cudaStream_t) { throw std::runtime_error("column-major: not implemented yet"); }
inline void colmajor_get_column(const Matrix *, Workspace *, u64, DeviceBuf<u64> *, u64 *) { throw std::runtime_error("column-major: not implemented yet"); }
}
