// rs_internal.h -- declarations shared by the .cu files of librsgpu.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace rsgpu {

// GF(2^8), primitive polynomial 0x11D (octal 0435: reference src/matrix.cu:49, src/cpu-rs.c:37).
constexpr unsigned kPoly = 0x11D;
constexpr int kExpSize = 512;   // exp[0..509] doubled so exp[log a + log b] needs no modulo; 510,511 = 0
constexpr int kLogSize = 256;

// Device-resident field tables: d_exp[kExpSize] then d_log[kLogSize], contiguous.
struct GfTables {
    uint8_t *d_exp = nullptr;
    uint8_t *d_log = nullptr;
};

// One group of output rows that a single multiply launch produces.  The product table holds, per
// input fragment j and nibble position (lo, hi), 16 entries of W bytes: entry n = the products of n
// (or n<<4) with the coefficients of up to W rows, one byte per row.
struct RowGroup {
    int row0 = 0;       // first output row of this group
    int nrows = 0;      // 1..8
    int width = 4;      // bytes per table entry: 4 (<= 4 rows) or 8 (<= 8 rows)
    bool all_ones = false;   // nrows == 1 and every coefficient is 1: the row is the plain XOR of the inputs (no lookups)
    int xor_row = -1;   // >= 0: this group's launch ALSO writes output row xor_row (= row0 - 1), whose coefficients
                        // are all 1, as the plain XOR of the inputs (the peeled row has no table entries)
    uint32_t *d_tab = nullptr;   // points into Plan::d_tabs
    size_t tab_words = 0;        // kpad * 2 * 16 * width / 4
};

constexpr int kMaxGroups = 32;   // 255 rows / 8

// A coding matrix (rows x k) ready to be applied: device copy of the matrix + product tables.
struct Plan {
    int rows = 0;
    int k = 0;
    int kpad = 0;                 // k rounded up to a multiple of 8 (zero tables behind k)
    int ngroups = 0;
    RowGroup grp[kMaxGroups];
    uint8_t *d_matrix = nullptr;  // rows x k
    uint32_t *d_tabs = nullptr;   // all groups back to back
    size_t tabs_bytes = 0;
    bool valid = false;
    bool peeled = false;          // row 0 is all ones and is produced by grp[0] as a XOR row (encode plans, p % 8 == 5)
};

struct Tuning {
    int variant = 0;      // 0 = default (see gf_matmul_sm100.cu)
    int threads = 0;      // 0 = default
    int max_blocks = 0;   // 0 = persistent grid from SM count
};

// ---- gf_tables.cu ---------------------------------------------------------------------------
cudaError_t launch_build_gf_tables(uint8_t *d_exp, uint8_t *d_log, cudaStream_t s);
cudaError_t launch_gen_total_matrix(uint8_t *d_total, int k, int p, const GfTables &gf, cudaStream_t s);
cudaError_t launch_gen_total_matrix_cauchy(uint8_t *d_total, int k, int p, const GfTables &gf, cudaStream_t s);
cudaError_t launch_rows_times_matrix(const uint8_t *d_A, const int *d_sel, int rows, const uint8_t *d_B, int k,
                                     uint8_t *d_C, const GfTables &gf, cudaStream_t s);
cudaError_t launch_gather_rows(const uint8_t *d_total, int k, const int *d_rows, int nrows,
                               uint8_t *d_out, cudaStream_t s);
cudaError_t launch_build_product_tables(const uint8_t *d_matrix, int k, int kpad, int row0, int nrows,
                                        int width, uint32_t *d_tab, const GfTables &gf, cudaStream_t s);

// ---- gj_inverse.cu --------------------------------------------------------------------------
// d_flag[0] = 0 ok, 1 singular.  d_in is not modified.  One block does the whole inversion.
// gather + invert + tables of the full and the erased-rows plan in one launch (rsgpu_set_survivors)
cudaError_t launch_decode_setup(const uint8_t *d_total, int k, int kpad, const int *survivors, const int *erased, int n_erased,
                                uint8_t *d_inv, uint32_t *d_full_tabs, uint8_t *d_erased_matrix, uint32_t *d_erased_tabs,
                                int *d_flag, const GfTables &gf, cudaStream_t s);
cudaError_t launch_gj_invert(const uint8_t *d_in, uint8_t *d_inv, int n, int *d_flag,
                             const GfTables &gf, cudaStream_t s);

// ---- gf_matmul_sm100.cu ---------------------------------------------------------------------
// Memo of launch set-up per (kernel, block size, dynamic shared memory): the opt-in shared-memory attribute
// has been raised and the resident blocks per SM are known, so a repeat launch costs no runtime queries
// (they are ~3-5 us per call, as much as a whole kernel on a 1 MiB stripe).  One memo per context (= per
// device and host thread); not thread-safe, like the context.
struct LaunchMemo {
    static constexpr int kSlots = 32;
    const void *fn[kSlots];
    int threads[kSlots];
    int smem[kSlots];
    int occ[kSlots];
    int n = 0;
};

struct DeviceInfo {
    int sm_count = 0;
    int max_smem_optin = 0;
    LaunchMemo *memo = nullptr;   // may be null: then every launch queries
};
// Raises cudaFuncAttributeMaxDynamicSharedMemorySize if needed and returns the resident blocks per SM.
cudaError_t prepare_launch(const DeviceInfo &dev, const void *fn, int threads, size_t smem, int *occ);
// out rows [g.row0, g.row0+g.nrows) = matrix rows * in, bytes [0, len).  *launches += kernels launched.
cudaError_t launch_matmul_group(const RowGroup &g, int k, int kpad,
                                const uint8_t *in, size_t in_pitch, uint8_t *out, size_t out_pitch,
                                size_t len, const Tuning &tune, const DeviceInfo &dev,
                                cudaStream_t s, long long *launches);


// ---- gf_matmul_fat_sm100.cu -------------------------------------------------------------------
// Bank-replicated byte-table kernel (one lookup per input byte).  cudaErrorNotSupported = no instance
// for this k: use the split-nibble kernel.  Handles the nvec 16-byte columns only.
cudaError_t launch_matmul_fat(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, uint8_t *out_g,
                              size_t out_pitch, size_t nvec, const DeviceInfo &dev, int max_blocks, int threads,
                              cudaStream_t s);


// ---- gf_matmul_fat_ua_sm100.cu -------------------------------------------------------------------
// The same kernel for rows of ANY alignment (aligned vector accesses, realigned in registers).  Codes the
// columns [*vec_lo, *vec_hi) of every row -- more precisely a superset-safe inner range: the caller codes
// [0, *vec_lo) and [*vec_hi, len) with the byte kernel.  cudaErrorNotSupported: no instance for k / too short.
cudaError_t launch_matmul_fat_ua(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, uint8_t *out_g,
                                 size_t out_pitch, size_t len, const DeviceInfo &dev, int max_blocks, cudaStream_t s,
                                 size_t *vec_lo, size_t *vec_hi);

// ---- gf_matmul_fat_mp_sm100.cu -------------------------------------------------------------------
// Byte tables in passes of 14 fragments, partial parities kept in registers (k >= 25).  Codes the first *done_vec
// 16-byte columns (whole block steps); cudaErrorNotSupported when the stripe is shorter than one step.
cudaError_t launch_matmul_fat_mp(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, uint8_t *out_g,
                                 size_t out_pitch, size_t nvec, const DeviceInfo &dev, int max_blocks, cudaStream_t s,
                                 size_t *done_vec);

// ---- gf_matmul_fat_batch_sm100.cu ---------------------------------------------------------------
// A batch of n_stripes stripes in ONE launch: fragment j of stripe s at in + s*in_stripe + j*in_pitch.
// cudaErrorNotSupported when there is no instance for k (the caller then launches per stripe).
cudaError_t launch_matmul_fat_batch(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, size_t in_stripe,
                                    uint8_t *out_g, size_t out_pitch, size_t out_stripe, size_t nvec, size_t n_stripes,
                                    const DeviceInfo &dev, int max_blocks, cudaStream_t s);

// last (len % 16) bytes of every stripe of a batch, one launch (gf_matmul_sm100.cu)
cudaError_t launch_matmul_bytes_batch(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, size_t in_stripe,
                                      uint8_t *out_g, size_t out_pitch, size_t out_stripe, size_t c0, unsigned rem,
                                      size_t n_stripes, const DeviceInfo &dev, cudaStream_t s);

// ---- gf_matmul_fat_cluster_sm100.cu -------------------------------------------------------------
// k = 17..32, at most 4 output rows: the fragments split over the two CTAs of a cluster, partial parities
// exchanged through distributed shared memory.  cudaErrorNotSupported otherwise.
cudaError_t launch_matmul_fat_cluster(const RowGroup &g, int k, const uint8_t *in, size_t in_pitch, uint8_t *out_g,
                                      size_t out_pitch, size_t nvec, const DeviceInfo &dev, int max_blocks, cudaStream_t s);

// ---- gf_matmul_rows_sm100.cu --------------------------------------------------------------------
// One destination pointer per output row (rows may be peer-GPU memory: the stores go over NVLink).
cudaError_t launch_matmul_rows(const RowGroup &g, int k, int kpad, const uint8_t *in, size_t in_pitch,
                               uint8_t *const *out_rows, size_t len, const DeviceInfo &dev, cudaStream_t s,
                               long long *launches);

// ---- gf_matmul_tma_sm100.cu ---------------------------------------------------------------------
// Split-nibble kernel fed by TMA bulk copies through a shared-memory stage ring (tuning variant 7).
// cudaErrorNotSupported when k has no divisor among the compiled batch sizes.
cudaError_t launch_matmul_tma(const RowGroup &g, int k, int kpad, const uint8_t *in, size_t in_pitch, uint8_t *out_g,
                              size_t out_pitch, size_t nvec, const DeviceInfo &dev, int max_blocks, cudaStream_t s);

}  // namespace rsgpu
