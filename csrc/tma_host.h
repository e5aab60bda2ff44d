// Host-side TMA descriptor helper shared by the kernels that feed tcgen05 through TMA (defined in gemm_sm100.cu).
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace eb {

// 2-D bf16 tensor map over a row-major [rows][ld] array whose logical extent is [rows][cols];
// box = [box_rows][box_cols], 128-byte swizzle, out-of-bounds elements read as zero.
int make_map(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
             uint32_t box_cols);

}  // namespace eb
