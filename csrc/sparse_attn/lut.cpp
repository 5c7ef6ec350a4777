// Host-side look-up-table builders for the block-sparse kernels + sdd_segment.
//   dsb_bs_build_luts : layout [H, nb, nb] (0/1) -> blocks (h, i, j), row CSR (col, blk), column CSC (row, blk)
//   dsb_sdd_segment   : greedy packing of a layout into max_width x max_width super-blocks, OpenMP over heads
// Same role as reference csrc/sparse_attention/utils.cpp:14-120 and the Python LUT code in
// deepspeed/ops/sparse_attention/matmul.py:98-331, softmax.py:22-41 (formats are ours).
#include <omp.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <vector>

extern "C" {

long dsb_bs_count_nnz(const int64_t* layout, int H, int nb) {
    long nnz = 0;
    const long total = (long)H * nb * nb;
    for (long i = 0; i < total; ++i) nnz += layout[i] != 0;
    return nnz;
}

// blocks: int32 [nnz, 4]; row_ptr/col_ptr: int32 [H*nb + 1]; row_idx/col_idx: int32 [nnz, 2]
int dsb_bs_build_luts(const int64_t* layout, int H, int nb, int32_t* blocks, int32_t* row_ptr, int32_t* row_idx,
                      int32_t* col_ptr, int32_t* col_idx) {
    long blk = 0;
    row_ptr[0] = 0;
    std::vector<int32_t> block_id((size_t)H * nb * nb, -1);
    for (int h = 0; h < H; ++h) {
        for (int i = 0; i < nb; ++i) {
            for (int j = 0; j < nb; ++j) {
                const long off = ((long)h * nb + i) * nb + j;
                if (layout[off] != 0) {
                    blocks[blk * 4 + 0] = h;
                    blocks[blk * 4 + 1] = i;
                    blocks[blk * 4 + 2] = j;
                    blocks[blk * 4 + 3] = 0;
                    row_idx[blk * 2 + 0] = j;
                    row_idx[blk * 2 + 1] = (int32_t)blk;
                    block_id[off] = (int32_t)blk;
                    ++blk;
                }
            }
            row_ptr[h * nb + i + 1] = (int32_t)blk;
        }
    }
    long c = 0;
    col_ptr[0] = 0;
    for (int h = 0; h < H; ++h) {
        for (int j = 0; j < nb; ++j) {
            for (int i = 0; i < nb; ++i) {
                const int32_t id = block_id[((long)h * nb + i) * nb + j];
                if (id >= 0) {
                    col_idx[c * 2 + 0] = i;
                    col_idx[c * 2 + 1] = id;
                    ++c;
                }
            }
            col_ptr[h * nb + j + 1] = (int32_t)c;
        }
    }
    return 0;
}

// Greedy segmentation: scan each head's layout for the largest square of ones (side <= max_width, power of two
// steps max_width, max_width/2, ..., 1) anchored at each still-unassigned block; every block gets the id of the
// super-block that covers it. out_id: int32 [H, nb, nb] (-1 for zeros); returns the number of super-blocks of each
// width in counts[log2(max_width)+1] (index = log2(width)).
int dsb_sdd_segment(const int64_t* layout, int H, int nb, int max_width, int32_t* out_id, int32_t* out_width,
                    int64_t* counts) {
    int levels = 0;
    for (int w = max_width; w >= 1; w >>= 1) ++levels;
    std::vector<std::vector<int64_t>> local(H, std::vector<int64_t>(levels, 0));
#pragma omp parallel for schedule(dynamic)
    for (int h = 0; h < H; ++h) {
        const int64_t* L = layout + (long)h * nb * nb;
        int32_t* id = out_id + (long)h * nb * nb;
        int32_t* wd = out_width + (long)h * nb * nb;
        std::fill(id, id + (long)nb * nb, -1);
        std::fill(wd, wd + (long)nb * nb, 0);
        int32_t next = 0;
        int level = levels - 1;
        for (int w = max_width; w >= 1; w >>= 1, --level) {
            for (int i = 0; i + w <= nb; i += w) {
                for (int j = 0; j + w <= nb; j += w) {
                    bool ok = true;
                    for (int a = 0; a < w && ok; ++a)
                        for (int b = 0; b < w; ++b)
                            if (L[(long)(i + a) * nb + j + b] == 0 || id[(long)(i + a) * nb + j + b] >= 0) { ok = false; break; }
                    if (!ok) continue;
                    for (int a = 0; a < w; ++a)
                        for (int b = 0; b < w; ++b) {
                            id[(long)(i + a) * nb + j + b] = next;
                            wd[(long)(i + a) * nb + j + b] = w;
                        }
                    ++next;
                    local[h][level]++;
                }
            }
        }
    }
    for (int l = 0; l < levels; ++l) {
        counts[l] = 0;
        for (int h = 0; h < H; ++h) counts[l] += local[h][l];
    }
    return 0;
}

}  // extern "C"
