// One-off layout kernels: scatter of reference-ordered (P|mn)/B columns into the tile-compacted HBM layout, and the
// deterministic synthetic shard used by benchmarks.
#include "dmma_common.cuh"
#include "jcdf_internal.h"

namespace jcdf {

// T_dev: (q_len x col_count) column-major window of the reference layout (rows = auxiliary functions of the shard,
// columns = kept (q,p) pairs in sparse_pq_index_map order, SchwarzScreening.jl:72-81).  dest_of_col maps a column to
// its offset inside a packed row (or -1 for the mirror copies that the q >= p tile storage does not keep).
// Reads are coalesced along Q; writes are 8-byte scatters (one-off setup).
__global__ void k_scatter_pairs(const double* __restrict__ T_dev, int64_t ld, int64_t col_begin, int64_t col_count,
                                int q_len, const int64_t* __restrict__ dest_of_col, double* __restrict__ Bp,
                                int64_t row_len) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= q_len) return;
    for (int64_t c = blockIdx.y; c < col_count; c += gridDim.y) {
        const int64_t dst = dest_of_col[col_begin + c];
        if (dst < 0) continue;
        Bp[(int64_t)q * row_len + dst] = T_dev[q + ld * c];
    }
}

void launch_scatter_pairs(const Layout& L, Shard& s, const double* T_dev, int64_t ld, int64_t col_begin,
                          int64_t col_count) {
    if (col_count <= 0 || s.q_len <= 0) return;
    dim3 block(128);
    dim3 grid((unsigned)((s.q_len + 127) / 128), (unsigned)std::min<int64_t>(col_count, 32768));
    k_scatter_pairs<<<grid, block, 0, s.stream>>>(T_dev, ld, col_begin, col_count, (int)s.q_len, s.d_dest_of_col, s.Bp,
                                                  L.row_len);
}

__global__ void k_set_identity(double* __restrict__ A, int64_t n) {
    const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (idx < n * n) A[idx] = (idx / n == idx % n) ? 1.0 : 0.0;
}
void launch_set_identity(double* A, int64_t n, cudaStream_t st) {
    k_set_identity<<<(unsigned)((n * n + 255) / 256), 256, 0, st>>>(A, n);
}

// inverse of k_scatter_pairs for introspection (jcdf_get_B): T_dev[q + ld * c] = Bp[q][src_of_col[col_begin + c]]
__global__ void k_gather_pairs(double* __restrict__ T_dev, int64_t ld, int64_t col_begin, int64_t col_count, int q_len,
                               const int64_t* __restrict__ src_of_col, const double* __restrict__ Bp, int64_t row_len) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= q_len) return;
    for (int64_t c = blockIdx.y; c < col_count; c += gridDim.y)
        T_dev[q + ld * c] = Bp[(int64_t)q * row_len + src_of_col[col_begin + c]];
}

void launch_gather_pairs(const Layout& L, Shard& s, double* T_dev, int64_t ld, int64_t col_begin, int64_t col_count,
                         const int64_t* d_src_of_col, int64_t row_begin, int64_t row_count) {
    if (col_count <= 0 || row_count <= 0) return;
    dim3 block(128);
    dim3 grid((unsigned)((row_count + 127) / 128), (unsigned)std::min<int64_t>(col_count, 32768));
    k_gather_pairs<<<grid, block, 0, s.stream>>>(T_dev, ld, col_begin, col_count, (int)row_count, d_src_of_col,
                                                 s.Bp + row_begin * L.row_len, L.row_len);
}

// ---------------------------------------------------------------------------------------------------------------------
// synthetic shard: integer-hash "Gaussian" (sum of four 16-bit uniforms) times a host-tabulated decay, so the host
// twin jcdf_synthetic_value reproduces every element bit for bit (one int->double conversion, one multiply).
// ---------------------------------------------------------------------------------------------------------------------
__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__host__ __device__ inline int synth_isum(uint64_t seed, int64_t q_aux, int64_t lo, int64_t hi) {
    uint64_t key = (((uint64_t)q_aux << 42) | ((uint64_t)lo << 21) | (uint64_t)hi);
    uint64_t h = splitmix64(key ^ splitmix64(seed));
    int s = (int)(h & 0xFFFF) + (int)((h >> 16) & 0xFFFF) + (int)((h >> 32) & 0xFFFF) + (int)((h >> 48) & 0xFFFF);
    return s - 2 * 65535;
}

__global__ void k_fill_synthetic(double* __restrict__ Bp, int64_t row_len, int q_len, int64_t q_begin,
                                 const int32_t* __restrict__ tile_a, const int32_t* __restrict__ tile_b,
                                 const uint8_t* __restrict__ fn_mask, const double* __restrict__ cdec, int N,
                                 uint64_t seed, int64_t n_tiles, const int32_t* __restrict__ inv) {
    const int c = threadIdx.x & 31;
    const int r0 = threadIdx.x >> 5;
    for (int64_t blk = blockIdx.x; blk < (int64_t)q_len * n_tiles; blk += gridDim.x) {
        const int64_t row = blk / n_tiles;
        const int tile = (int)(blk % n_tiles);
        const int a = tile_a[tile], b = tile_b[tile];
        const int qi = b * 32 + c;
        for (int r = r0; r < 32; r += blockDim.x >> 5) {
            const int pi = a * 32 + r;
            double v = 0.0;
            const int p = pi < N ? inv[pi] : pi, q = qi < N ? inv[qi] : qi;  // the caller's function indices
            if (pi < N && qi < N && (fn_mask == nullptr || fn_mask[p + (int64_t)N * q])) {
                const int lo = min(p, q), hi = max(p, q);
                v = (double)synth_isum(seed, q_begin + row, lo, hi) * cdec[hi - lo];
            }
            Bp[row * row_len + (int64_t)tile * TILE_ELEMS + r * 32 + swz(r, c)] = v;
        }
    }
}

static void synth_table(std::vector<double>& cdec, int64_t N, int64_t Q_total, double decay_length) {
    cdec.resize((size_t)N);
    const double sigma = 65536.0 / 1.7320508075688772;  // std of the sum of four 16-bit uniforms
    for (int64_t d = 0; d < N; d++) {
        double dec = decay_length > 0 ? exp(-(double)d / decay_length) : 1.0;
        cdec[(size_t)d] = dec / (sigma * sqrt((double)Q_total));
    }
}

void launch_fill_synthetic(const Layout& L, Shard& s, uint64_t seed, double decay_length) {
    std::vector<double> cdec;
    synth_table(cdec, L.N, L.Q, decay_length);
    double* d_cdec = nullptr;
    cudaMalloc(&d_cdec, sizeof(double) * cdec.size());
    cudaMemcpyAsync(d_cdec, cdec.data(), sizeof(double) * cdec.size(), cudaMemcpyHostToDevice, s.stream);
    int64_t blocks = (int64_t)s.q_len * L.tiles_kept;
    int grid = (int)std::min<int64_t>(blocks, 148 * 32);
    if (grid > 0)
        k_fill_synthetic<<<grid, 256, 0, s.stream>>>(s.Bp, L.row_len, (int)s.q_len, s.q_begin, s.d_tile_a, s.d_tile_b,
                                                     s.d_fn_mask, d_cdec, (int)L.N, seed, L.tiles_kept, s.d_inv);
    cudaStreamSynchronize(s.stream);
    cudaFree(d_cdec);
}

}  // namespace jcdf

namespace jcdf {
__global__ void k_checked_selftest(const double* base, int64_t elems, int64_t index) {
    JC_CHECK(index >= 0 && index < elems, "self-test: index outside its buffer (expected)");
    if (index >= 0 && index < elems && base[index] == 12345.0) printf("\n");
}
}  // namespace jcdf

// 1 when this library was built with JCDF_CHECKED (make CHECKED=1), else 0.  With fire != 0 a one-thread kernel is launched whose
// JC_CHECK must fail: the call then returns the CUDA error code the trap produces (non-zero; the context is dead afterwards —
// call it last, in a process of its own).  Proves that the bounds checks of the checked build are live.
extern "C" int32_t jcdf_checked_selftest(int32_t fire) {
#ifdef JCDF_CHECKED
    if (!fire) return 1;
    double* d = nullptr;
    if (cudaMalloc(&d, 64) != cudaSuccess) return -2;
    jcdf::k_checked_selftest<<<1, 1>>>(d, 8, 8);
    const cudaError_t e = cudaDeviceSynchronize();
    return e == cudaSuccess ? -3 : (int32_t)e;
#else
    (void)fire;
    return 0;
#endif
}

extern "C" double jcdf_synthetic_value(uint64_t seed, int64_t Q_total, int64_t q_aux, int64_t p, int64_t q,
                                       double decay_length) {
    const int64_t lo = p < q ? p : q, hi = p < q ? q : p;
    const double sigma = 65536.0 / 1.7320508075688772;
    double dec = decay_length > 0 ? exp(-(double)(hi - lo) / decay_length) : 1.0;
    double cd = dec / (sigma * sqrt((double)Q_total));
    return (double)jcdf::synth_isum(seed, q_aux, lo, hi) * cd;
}
