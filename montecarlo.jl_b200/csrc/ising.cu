// Classical 2D Ising checkerboard Metropolis, bit-packed (1 bit/spin, two colour planes), with a
// Philox4x32-10 counter RNG and integer acceptance thresholds.
//
// Replaces  sweep(mc::MC)                 src/flavors/MC/MC.jl:316-332
//           propose_local / accept_local! src/models/Ising/IsingModel.jl:83-101
//           energy, sum(conf)             src/models/Ising/IsingModel.jl:177-185, Ising/measurements.jl:72-78
//
// Layout: site (x,y), colour c = (x+y)&1. Plane c, row y holds the L/2 sites of that colour in
// order of xh = x>>1 (x = 2*xh + ((y+c)&1)); 64 consecutive xh per uint64 word, W words per row:
//   plane[replica][c][y][w], bit b of word w  <->  xh = 64*w + b.
// Neighbours (all in the other plane): rows y+-1 at the same (w,b); in row y at xh and xh+1
// (if (y+c) is odd) or xh-1 and xh (if even).
// RNG: key = (seed, replica), counter = (sweep, colour, (y*W + w)*16 + b/4, 0), word b%4.
#include "../../include/mcb200.h"
#include "common.cuh"
#include <math.h>

using namespace mcb;

struct mcb_ising {
    int L = 0, R = 0, W = 0, nbits = 0; // nbits = valid bits in a word (64 or L/2 when L/2 < 64)
    double beta = 0;
    uint32_t thr4 = 0, thr8 = 0;
    uint64_t seed = 0;
    uint32_t replica_offset = 0;
    int device = 0;
    cudaStream_t st = nullptr;
    uint64_t* planes = nullptr;   // [R][2][L][W]
    int8_t* spins_dev = nullptr;  // [R][L][L] staging for set/get
    unsigned long long* red = nullptr; // [R][2] : aligned-bond count, up-spin count
    int64_t sweeps_done = 0;
    int64_t launches = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0.f;
    bool spins_set = false;
};

namespace {

struct IsingGeom { int L, R, W, nbits; };

__device__ __forceinline__ uint64_t word_mask(int nbits) { return nbits == 64 ? ~0ull : ((1ull << nbits) - 1ull); }

// other-plane word shifted so that bit b holds the neighbour at xh+1 (dir=+1) or xh-1 (dir=-1)
__device__ __forceinline__ uint64_t shifted_row(const uint64_t* row, int w, int dir, const IsingGeom& g)
{
    const uint64_t o = row[w];
    if (g.W == 1) {
        const uint64_t m = word_mask(g.nbits);
        return dir > 0 ? (((o >> 1) | (o << (g.nbits - 1))) & m) : (((o << 1) | (o >> (g.nbits - 1))) & m);
    }
    if (dir > 0) return (o >> 1) | (row[w + 1 == g.W ? 0 : w + 1] << 63);
    return (o << 1) | (row[w == 0 ? g.W - 1 : w - 1] >> 63);
}

__global__ void __launch_bounds__(256) ising_halfsweep_kernel(uint64_t* planes, IsingGeom g, int colour,
                                                              uint32_t sweep, uint32_t key0,
                                                              uint32_t replica_offset, uint32_t thr4,
                                                              uint32_t thr8)
{
    const size_t words_per_plane = (size_t)g.L * g.W;
    const size_t total = (size_t)g.R * words_per_plane;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int rep = (int)(idx / words_per_plane);
        const size_t rem = idx % words_per_plane;
        const int y = (int)(rem / g.W), w = (int)(rem % g.W);
        uint64_t* own_plane = planes + ((size_t)rep * 2 + colour) * words_per_plane;
        const uint64_t* oth = planes + ((size_t)rep * 2 + (colour ^ 1)) * words_per_plane;
        const uint64_t own = own_plane[rem];
        const uint64_t* row = oth + (size_t)y * g.W;
        const uint64_t n_same = row[w];
        const uint64_t n_shift = shifted_row(row, w, ((y + colour) & 1) ? +1 : -1, g);
        const uint64_t n_yp = oth[(size_t)(y + 1 == g.L ? 0 : y + 1) * g.W + w];
        const uint64_t n_ym = oth[(size_t)(y == 0 ? g.L - 1 : y - 1) * g.W + w];
        // aligned-neighbour count per bit (0..4) with bit-sliced adders
        const uint64_t a1 = ~(n_same ^ own), a2 = ~(n_shift ^ own), a3 = ~(n_yp ^ own), a4 = ~(n_ym ^ own);
        const uint64_t s0 = a1 ^ a2, c0 = a1 & a2, s1 = a3 ^ a4, c1 = a3 & a4;
        const uint64_t b0 = s0 ^ s1, carry = s0 & s1;
        const uint64_t b1 = c0 ^ c1 ^ carry, b2 = c0 & c1;
        const uint64_t eq3 = b1 & b0 & ~b2, eq4 = b2;   // dE = +4, +8
        const uint64_t le2 = ~(eq3 | eq4);              // dE <= 0 : always accepted
        // acceptance words: r < thr(dE)
        uint64_t m4 = 0, m8 = 0;
        const uint32_t gbase = (uint32_t)(((size_t)y * g.W + w) * 16);
        const uint32_t k1 = replica_offset + (uint32_t)rep;
#pragma unroll 4
        for (int q = 0; q < 16; ++q) {
            uint32_t r[4];
            philox4x32_10(sweep, (uint32_t)colour, gbase + q, 0u, key0, k1, r);
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                m4 |= (uint64_t)(r[t] < thr4) << (4 * q + t);
                m8 |= (uint64_t)(r[t] < thr8) << (4 * q + t);
            }
        }
        const uint64_t acc = (le2 | (eq3 & m4) | (eq4 & m8)) & word_mask(g.nbits);
        own_plane[rem] = own ^ acc;
    }
}

__global__ void ising_pack_kernel(const int8_t* spins, uint64_t* planes, IsingGeom g)
{
    const size_t words_per_plane = (size_t)g.L * g.W;
    const size_t total = (size_t)g.R * 2 * words_per_plane;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int rep = (int)(idx / (2 * words_per_plane));
        const size_t rem2 = idx % (2 * words_per_plane);
        const int colour = (int)(rem2 / words_per_plane);
        const size_t rem = rem2 % words_per_plane;
        const int y = (int)(rem / g.W), w = (int)(rem % g.W);
        const int8_t* s = spins + (size_t)rep * g.L * g.L + (size_t)y * g.L;
        uint64_t word = 0;
        for (int b = 0; b < g.nbits; ++b) {
            const int x = 2 * (64 * w + b) + ((y + colour) & 1);
            word |= (uint64_t)(s[x] > 0) << b;
        }
        planes[idx] = word;
    }
}

__global__ void ising_unpack_kernel(int8_t* spins, const uint64_t* planes, IsingGeom g)
{
    const size_t sites = (size_t)g.L * g.L;
    const size_t total = (size_t)g.R * sites;
    const size_t words_per_plane = (size_t)g.L * g.W;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int rep = (int)(idx / sites);
        const size_t site = idx % sites;
        const int x = (int)(site % g.L), y = (int)(site / g.L);
        const int colour = (x + y) & 1, xh = x >> 1;
        const uint64_t word = planes[((size_t)rep * 2 + colour) * words_per_plane + (size_t)y * g.W + (xh >> 6)];
        spins[idx] = ((word >> (xh & 63)) & 1ull) ? 1 : -1;
    }
}

// Spins in the recorder's wire format, BitArray(conf .== 1): per replica ceil(L*L/64) little-endian 64-bit chunks,
// bit s = site x + y*L. One thread per chunk (not a hot path: one pass over L*L bits per call).
__global__ void ising_planes_to_bits_kernel(uint64_t* bits, const uint64_t* planes, IsingGeom g, int chunks)
{
    const size_t total = (size_t)g.R * chunks, sites = (size_t)g.L * g.L;
    const size_t words_per_plane = (size_t)g.L * g.W;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int rep = (int)(idx / chunks);
        const size_t s0 = (idx % chunks) * 64;
        uint64_t out = 0;
        for (int b = 0; b < 64 && s0 + b < sites; ++b) {
            const size_t site = s0 + b;
            const int x = (int)(site % g.L), y = (int)(site / g.L);
            const int colour = (x + y) & 1, xh = x >> 1;
            const uint64_t word = planes[((size_t)rep * 2 + colour) * words_per_plane + (size_t)y * g.W + (xh >> 6)];
            out |= ((word >> (xh & 63)) & 1ull) << b;
        }
        bits[idx] = out;
    }
}

__global__ void ising_bits_to_planes_kernel(const uint64_t* bits, uint64_t* planes, IsingGeom g, int chunks)
{
    const size_t words_per_plane = (size_t)g.L * g.W;
    const size_t total = (size_t)g.R * 2 * words_per_plane;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int rep = (int)(idx / (2 * words_per_plane));
        const size_t rem2 = idx % (2 * words_per_plane);
        const int colour = (int)(rem2 / words_per_plane);
        const size_t rem = rem2 % words_per_plane;
        const int y = (int)(rem / g.W), w = (int)(rem % g.W);
        const uint64_t* src = bits + (size_t)rep * chunks;
        uint64_t word = 0;
        for (int b = 0; b < g.nbits; ++b) {
            const size_t site = (size_t)(2 * (64 * w + b) + ((y + colour) & 1)) + (size_t)y * g.L;
            word |= ((src[site >> 6] >> (site & 63)) & 1ull) << b;
        }
        planes[idx] = word;
    }
}

// red[rep][0] += #aligned (site, up) + #aligned (site, right) bonds ; red[rep][1] += #up spins
__global__ void ising_measure_kernel(const uint64_t* planes, IsingGeom g, unsigned long long* red)
{
    const size_t words_per_plane = (size_t)g.L * g.W;
    const size_t per_rep = 2 * words_per_plane;
    const int rep = blockIdx.y;
    unsigned long long aligned = 0, ups = 0;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < per_rep;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int colour = (int)(idx / words_per_plane);
        const size_t rem = idx % words_per_plane;
        const int y = (int)(rem / g.W), w = (int)(rem % g.W);
        const uint64_t own = planes[(size_t)rep * per_rep + idx];
        const uint64_t* oth = planes + ((size_t)rep * 2 + (colour ^ 1)) * words_per_plane;
        const uint64_t* row = oth + (size_t)y * g.W;
        // up = (x+1, y): other-plane position xh+1 if (y+colour) odd, xh if even
        const uint64_t n_up = ((y + colour) & 1) ? shifted_row(row, w, +1, g) : row[w];
        const uint64_t n_right = oth[(size_t)(y + 1 == g.L ? 0 : y + 1) * g.W + w];
        const uint64_t m = word_mask(g.nbits);
        aligned += __popcll(~(own ^ n_up) & m) + __popcll(~(own ^ n_right) & m);
        ups += __popcll(own & m);
    }
    for (int o = 16; o > 0; o >>= 1) {
        aligned += __shfl_xor_sync(0xffffffffu, aligned, o);
        ups += __shfl_xor_sync(0xffffffffu, ups, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&red[2 * rep], aligned);
        atomicAdd(&red[2 * rep + 1], ups);
    }
}

IsingGeom geom(const mcb_ising* h) { return IsingGeom{h->L, h->R, h->W, h->nbits}; }

int check_device(int dev)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        set_error("no CUDA device: libmcb200 has no CPU fallback");
        return MCB_ERR_CUDA;
    }
    MCB_CUDA(cudaSetDevice(dev));
    return 0;
}

} // namespace

extern "C" {

int mcb_ising_create(int32_t L, double beta, int32_t n_replicas, uint64_t seed, uint32_t replica_offset,
                     int32_t device_id, mcb_ising** out)
{
    if (!out || L < 2 || (L & 1) || n_replicas <= 0) { set_error("L must be even and >= 2, replicas > 0"); return MCB_ERR_ARG; }
    if (L / 2 > 64 && (L / 2) % 64 != 0) { set_error("L/2 must be <= 64 or a multiple of 64"); return MCB_ERR_ARG; }
    if (n_replicas > 65535) { set_error("at most 65535 replicas per handle"); return MCB_ERR_ARG; }
    if (!(beta >= 0.0) || !isfinite(beta)) { set_error("beta must be finite and >= 0"); return MCB_ERR_ARG; }
    const int rc = check_device(device_id);
    if (rc) return rc;
    mcb_ising* h = new mcb_ising();
    h->L = L; h->R = n_replicas; h->beta = beta; h->seed = seed; h->replica_offset = replica_offset;
    h->device = device_id;
    h->W = (L / 2 + 63) / 64;
    h->nbits = (L / 2 >= 64) ? 64 : L / 2;
    // thresholds in double on the host, compared as integers on the device (bit-exact by construction)
    // (saturated: at beta = 0 the product is 2^32, which does not fit; r < 2^32 - 1 then accepts all but one word)
    h->thr4 = (uint32_t)fmin(floor(exp(-4.0 * beta) * 4294967296.0), 4294967295.0);
    h->thr8 = (uint32_t)fmin(floor(exp(-8.0 * beta) * 4294967296.0), 4294967295.0);
    MCB_CUDA(cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking));
    MCB_CUDA(cudaMalloc(&h->planes, (size_t)h->R * 2 * h->L * h->W * sizeof(uint64_t)));
    MCB_CUDA(cudaMalloc(&h->spins_dev, (size_t)h->R * h->L * h->L));
    MCB_CUDA(cudaMalloc(&h->red, (size_t)h->R * 2 * sizeof(unsigned long long)));
    MCB_CUDA(cudaEventCreate(&h->ev0));
    MCB_CUDA(cudaEventCreate(&h->ev1));
    *out = h;
    return MCB_OK;
}

int mcb_ising_destroy(mcb_ising* h)
{
    if (!h) return MCB_OK;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->st);
    cudaFree(h->planes); cudaFree(h->spins_dev); cudaFree(h->red);
    cudaEventDestroy(h->ev0); cudaEventDestroy(h->ev1);
    cudaStreamDestroy(h->st);
    delete h;
    return MCB_OK;
}

int mcb_ising_set_spins(mcb_ising* h, const int8_t* spins)
{
    if (!h || !spins) { set_error("null argument"); return MCB_ERR_ARG; }
    MCB_CUDA(cudaSetDevice(h->device));
    MCB_CUDA(cudaMemcpyAsync(h->spins_dev, spins, (size_t)h->R * h->L * h->L, cudaMemcpyHostToDevice, h->st));
    ising_pack_kernel<<<592, 256, 0, h->st>>>(h->spins_dev, h->planes, geom(h));
    MCB_CUDA(cudaGetLastError());
    MCB_CUDA(cudaStreamSynchronize(h->st));
    h->spins_set = true;
    return MCB_OK;
}

int mcb_ising_get_spins(mcb_ising* h, int8_t* spins)
{
    if (!h || !spins) { set_error("null argument"); return MCB_ERR_ARG; }
    MCB_CUDA(cudaSetDevice(h->device));
    ising_unpack_kernel<<<1184, 256, 0, h->st>>>(h->spins_dev, h->planes, geom(h));
    MCB_CUDA(cudaGetLastError());
    MCB_CUDA(cudaMemcpyAsync(spins, h->spins_dev, (size_t)h->R * h->L * h->L, cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_ising_spin_chunks(mcb_ising* h) { return h ? (int)(((size_t)h->L * h->L + 63) / 64) : 0; }

int mcb_ising_set_spins_bits(mcb_ising* h, const uint64_t* chunks)
{
    if (!h || !chunks) { set_error("null argument"); return MCB_ERR_ARG; }
    MCB_CUDA(cudaSetDevice(h->device));
    const int nch = mcb_ising_spin_chunks(h);
    uint64_t* dev = reinterpret_cast<uint64_t*>(h->spins_dev);   // R*L*L bytes >= R*nch*8
    MCB_CUDA(cudaMemcpyAsync(dev, chunks, (size_t)h->R * nch * sizeof(uint64_t), cudaMemcpyHostToDevice, h->st));
    ising_bits_to_planes_kernel<<<592, 256, 0, h->st>>>(dev, h->planes, geom(h), nch);
    MCB_CUDA(cudaGetLastError());
    MCB_CUDA(cudaStreamSynchronize(h->st));
    h->spins_set = true;
    return MCB_OK;
}

int mcb_ising_get_spins_bits(mcb_ising* h, uint64_t* chunks)
{
    if (!h || !chunks) { set_error("null argument"); return MCB_ERR_ARG; }
    MCB_CUDA(cudaSetDevice(h->device));
    const int nch = mcb_ising_spin_chunks(h);
    uint64_t* dev = reinterpret_cast<uint64_t*>(h->spins_dev);
    ising_planes_to_bits_kernel<<<592, 256, 0, h->st>>>(dev, h->planes, geom(h), nch);
    MCB_CUDA(cudaGetLastError());
    MCB_CUDA(cudaMemcpyAsync(chunks, dev, (size_t)h->R * nch * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    return MCB_OK;
}

int mcb_ising_sweep(mcb_ising* h, int64_t n_sweeps)
{
    if (!h || n_sweeps < 0) { set_error("bad argument"); return MCB_ERR_ARG; }
    if (!h->spins_set) { set_error("set_spins first"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->device));
    const size_t total = (size_t)h->R * h->L * h->W;
    const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
    MCB_CUDA(cudaEventRecord(h->ev0, h->st));
    for (int64_t s = 0; s < n_sweeps; ++s) {
        for (int colour = 0; colour < 2; ++colour)
            ising_halfsweep_kernel<<<blocks, 256, 0, h->st>>>(h->planes, geom(h), colour,
                                                              (uint32_t)h->sweeps_done, seed_key0(h->seed),
                                                              h->replica_offset, h->thr4, h->thr8);
        h->sweeps_done += 1;
        h->launches += 2;
    }
    MCB_CUDA(cudaGetLastError());
    MCB_CUDA(cudaEventRecord(h->ev1, h->st));
    MCB_CUDA(cudaStreamSynchronize(h->st));
    MCB_CUDA(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
    return MCB_OK;
}

int mcb_ising_measure(mcb_ising* h, int64_t* E, int64_t* Msum)
{
    if (!h || !E || !Msum) { set_error("null argument"); return MCB_ERR_ARG; }
    if (!h->spins_set) { set_error("set_spins first"); return MCB_ERR_STATE; }
    MCB_CUDA(cudaSetDevice(h->device));
    MCB_CUDA(cudaMemsetAsync(h->red, 0, (size_t)h->R * 2 * sizeof(unsigned long long), h->st));
    const size_t per_rep = (size_t)2 * h->L * h->W;
    int bx = (int)((per_rep + 255) / 256);
    if (bx > 64) bx = 64;
    ising_measure_kernel<<<dim3(bx, h->R), 256, 0, h->st>>>(h->planes, geom(h), h->red);
    MCB_CUDA(cudaGetLastError());
    std::string dummy;
    unsigned long long* host = new unsigned long long[(size_t)h->R * 2];
    cudaError_t err = cudaMemcpyAsync(host, h->red, (size_t)h->R * 2 * sizeof(unsigned long long),
                                      cudaMemcpyDeviceToHost, h->st);
    if (err == cudaSuccess) err = cudaStreamSynchronize(h->st);
    if (err != cudaSuccess) { delete[] host; set_error(cudaGetErrorString(err)); return MCB_ERR_CUDA; }
    const int64_t sites = (int64_t)h->L * h->L;
    for (int r = 0; r < h->R; ++r) {
        // E = -sum_i s_i (s_up + s_right) = -(2 * aligned - 2 * sites)
        E[r] = -(2 * (int64_t)host[2 * r] - 2 * sites);
        Msum[r] = 2 * (int64_t)host[2 * r + 1] - sites;
    }
    delete[] host;
    return MCB_OK;
}

int64_t mcb_ising_sweep_count(mcb_ising* h) { return h ? h->sweeps_done : -1; }
double mcb_ising_last_elapsed_ms(mcb_ising* h) { return h ? (double)h->last_ms : -1.0; }
int64_t mcb_ising_launch_count(mcb_ising* h) { return h ? h->launches : -1; }

} // extern "C"
