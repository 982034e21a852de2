// Stand-alone check of the TMA usage pattern of k_cost (2-D u32 tiles, negative / OOB coordinates,
// two tensor maps on one mbarrier, double-buffered re-arming).
#include <cstdio>
#include <cstdint>
#include <vector>
#include "../stereovision_b200/csrc/svb_common.cuh"
#include "../stereovision_b200/csrc/core.cu"
__global__ void k(const __grid_constant__ CUtensorMap tm, int x, int y, int bw, int rows, uint32_t* out)
{
    extern __shared__ __align__(128) uint32_t sm[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x == 0) { mbar_expect_tx(&bar, bw * rows * 4); tma_load_2d(sm, &tm, x, y, &bar); }
    mbar_wait(&bar, 0);
    for (int i = threadIdx.x; i < bw * rows; i += blockDim.x) out[i] = sm[i];
}
// k_cost-like: per "row" two boxes (6 x BW from tmR, 1 x 256 from tmL) into buffer it&1, re-armed after a block barrier
__global__ void k2(const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmL, int xR, int xL, int nrows,
                   int BW, unsigned long long* sum)
{
    extern __shared__ __align__(128) uint32_t sm[];
    __shared__ __align__(8) uint64_t mbar[2];
    uint32_t* stage = sm;                 // [2][6][BW]
    uint32_t* lstage = sm + 2 * 6 * BW;   // [2][256]
    auto issue = [&](int v, int buf) {
        mbar_expect_tx(&mbar[buf], (6 * BW + 256) * 4);
        tma_load_2d(stage + buf * 6 * BW, &tmR, xR, v * 6, &mbar[buf]);
        tma_load_2d(lstage + buf * 256, &tmL, xL, v, &mbar[buf]);
    };
    if (threadIdx.x == 0) { mbar_init(&mbar[0], 1); mbar_init(&mbar[1], 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x == 0) { issue(0, 0); if (nrows > 1) issue(1, 1); }
    unsigned long long s = 0;
    for (int it = 0; it < nrows; ++it) {
        const int buf = it & 1;
        mbar_wait(&mbar[buf], (it >> 1) & 1);
        for (int i = threadIdx.x; i < 6 * BW; i += blockDim.x) s += stage[buf * 6 * BW + i];
        for (int i = threadIdx.x; i < 256; i += blockDim.x) s += lstage[buf * 256 + i];
        __syncthreads();
        if (threadIdx.x == 0 && it + 2 < nrows) issue(it + 2, buf);
    }
    atomicAdd(sum, s);
}
int main()
{
    const int W = 132, R = 360;
    std::vector<uint32_t> h(W * R);
    for (int i = 0; i < W * R; ++i) h[i] = (i / W) * 1000 + i % W + 1;
    uint32_t *d, *o; cudaMalloc(&d, h.size() * 4); cudaMalloc(&o, 4096 * 4);
    cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    struct { int bw, rows, x, y; } cases[] = {{64, 6, 0, 0}, {64, 6, -16, 6}, {64, 6, 100, 354}, {128, 1, 120, 3}, {64, 6, 14, 12}, {64, 6, 45, 12}, {64, 6, -3, 12}};
    for (auto c : cases) {
        CUtensorMap tm;
        if (svb_make_tmap_2d_u32(&tm, d, W, R, W * 4, c.bw, c.rows)) { printf("encode failed\n"); continue; }
        cudaMemset(o, 0xff, 4096 * 4);
        k<<<1, 128, c.bw * c.rows * 4>>>(tm, c.x, c.y, c.bw, c.rows, o);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<uint32_t> r(c.bw * c.rows);
        cudaMemcpy(r.data(), o, r.size() * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int j = 0; j < c.rows; ++j) for (int i = 0; i < c.bw; ++i) {
            int xx = c.x + i, yy = c.y + j;
            uint32_t want = (xx >= 0 && xx < W && yy >= 0 && yy < R) ? h[yy * W + xx] : 0;
            bad += r[j * c.bw + i] != want;
        }
        printf("box %dx%d at (%d,%d): %s, mismatches %d\n", c.bw, c.rows, c.x, c.y, cudaGetErrorString(e), bad);
        if (e != cudaSuccess) return 1;
    }
    // k_cost-like
    const int H = 60, Wl = 96 * 8;
    std::vector<uint32_t> hl(Wl * H, 1);
    uint32_t* dl; cudaMalloc(&dl, hl.size() * 4); cudaMemcpy(dl, hl.data(), hl.size() * 4, cudaMemcpyHostToDevice);
    unsigned long long* ds; cudaMalloc(&ds, 8); cudaMemset(ds, 0, 8);
    CUtensorMap tmR, tmL;
    svb_make_tmap_2d_u32(&tmR, d, W, R, W * 4, 64, 6);
    svb_make_tmap_2d_u32(&tmL, dl, Wl, H, Wl * 4, 256, 1);
    for (int xl : {120, 600, 900}) {
        k2<<<2, 512, (2 * 6 * 64 + 2 * 256) * 4>>>(tmR, tmL, -16, xl, 40, 64, ds);
        cudaError_t e = cudaDeviceSynchronize();
        unsigned long long hs; cudaMemcpy(&hs, ds, 8, cudaMemcpyDeviceToHost);
        printf("k2 xL=%d: %s sum %llu\n", xl, cudaGetErrorString(e), hs);
        if (e != cudaSuccess) return 1;
    }
    return 0;
}
