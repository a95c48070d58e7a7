// noalloc.cu -- amvpt_intersect_static: the ray tap of amvpt_intersect (closest-hit / shadow-ray BVH traversal, capi.cpp) for callers that
// must not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// amvpt_intersect stages the rays, the results and the traversal counters in device buffers it allocates per call.  Here that scratch is ONE
// file-scope device array of a fixed size (AMVPT_STATIC_MAX_RAYS rays x 44 B + 64 B of counters = 2.75 MiB): every GPU that loads the library
// holds it until exit.  The entry point calls no cudaMalloc / cudaFree / allocating library routine: cudaMemcpyAsync / cudaMemsetAsync into
// the array, one kernel, cudaMemcpyAsync out, one stream synchronise.
//
// Every call shares the array, and it keeps the previous call's contents: the counters are cleared at the start of every call (they are the only
// part that must start at zero; rays and results are overwritten for the n rays of the call before they are read).  Calls that overlap on
// the GPU (two host threads, or two streams) RACE on it -- callers serialise.
#include "../../include/amvpt.h"
#include "dpath.cuh"
#include "host.h"
#include <functional>
#include <mutex>

extern "C" int amvpt_guard_call(const std::function<int()> &fn);      // capi.cpp: exception -> error code + thread-local message
struct amvpt_scene { amvpt::Scene *s; };

namespace amvpt {
namespace {

constexpr uint32_t MAX_RAYS = AMVPT_STATIC_MAX_RAYS;
// layout of the scratch (byte offsets, each a multiple of 256): o[3n] | d[3n] | maxt[n] | prim[n] | tuv[3n] | counters[CTR_COUNT]
constexpr size_t OFF_O = 0, OFF_D = OFF_O + (size_t) MAX_RAYS * 12, OFF_M = OFF_D + (size_t) MAX_RAYS * 12, OFF_P = OFF_M + (size_t) MAX_RAYS * 4,
                 OFF_T = OFF_P + (size_t) MAX_RAYS * 4, OFF_C = OFF_T + (size_t) MAX_RAYS * 12, SCRATCH_BYTES = OFF_C + 8 * CTR_COUNT;
static_assert(OFF_D % 256 == 0 && OFF_M % 256 == 0 && OFF_P % 256 == 0 && OFF_T % 256 == 0 && OFF_C % 256 == 0, "scratch sections must stay 256-byte aligned");

// In host code this name is a host-side stand-in, NOT a device address: the kernel gets the address cudaGetSymbolAddress returns (below).
static __device__ __align__(256) unsigned char g_static_scratch[SCRATCH_BYTES];

// Same per-ray work as intersect_kernel (kernels.cu): the shared intersect<ANY, Q4> of dscene.cuh, the primitive id of the hit triangle,
// (t, u, v), and the node / triangle counters -- reading the rays from, and writing the results to, the scratch it is handed.
__global__ void intersect_static_kernel(const DScene s, unsigned char *scratch, uint32_t n, int any) {
    const float *o = reinterpret_cast<const float *>(scratch + OFF_O), *d = reinterpret_cast<const float *>(scratch + OFF_D),
                *maxt = reinterpret_cast<const float *>(scratch + OFF_M);
    uint32_t *prim = reinterpret_cast<uint32_t *>(scratch + OFF_P); float *tuv = reinterpret_cast<float *>(scratch + OFF_T);
    unsigned long long *counters = reinterpret_cast<unsigned long long *>(scratch + OFF_C);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    StatCtx st; st.closest = st.shadow = st.nodes = st.tris = st.vertices = 0;
    if (i < n) {
        Ray r; r.o = v3(o[3 * i], o[3 * i + 1], o[3 * i + 2]); r.d = v3(d[3 * i], d[3 * i + 1], d[3 * i + 2]); r.maxt = maxt[i];
        const bool q4 = (any & 2) != 0 && s.nodes_q4 != nullptr;
        Hit h = (any & 1) ? (q4 ? intersect<true, true>(s, r, &st.nodes, &st.tris) : intersect<true>(s, r, &st.nodes, &st.tris))
                          : (q4 ? intersect<false, true>(s, r, &st.nodes, &st.tris) : intersect<false>(s, r, &st.nodes, &st.tris));
        prim[i] = h.tri == AM_NO_HIT ? AM_NO_HIT : __float_as_uint(s.tris[3 * (size_t) h.tri].w);
        tuv[3 * i] = h.t; tuv[3 * i + 1] = h.u; tuv[3 * i + 2] = h.v;
        if (any & 1) st.shadow++; else st.closest++;
    }
    flush_stats(st, counters);
}

// Device address of the scratch, resolved once per device with that device current.
constexpr int MAX_DEVICES = 64;
std::mutex g_mutex; unsigned char *g_scratch_dev[MAX_DEVICES] = { nullptr };
unsigned char *scratch_on(int device) {
    if (device < 0 || device >= MAX_DEVICES) AM_THROW("amvpt_intersect_static: device %d out of range", device);
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_scratch_dev[device]) {
        AM_CUDA(cudaSetDevice(device));
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) AM_THROW("CUDA error: cudaGetSymbolAddress(static scratch) on device %d: %s", device, cudaGetErrorString(e));
        g_scratch_dev[device] = static_cast<unsigned char *>(p);
    }
    return g_scratch_dev[device];
}

} // namespace
} // namespace amvpt

extern "C" int amvpt_intersect_static(amvpt_scene *h, const float *o, const float *d, const float *maxt, uint32_t n, int any_hit,
                                      uint32_t *prim, float *tuv, uint64_t *stats2) {
    using namespace amvpt;
    return amvpt_guard_call([&]() -> int {
        if (!h || !h->s) AM_THROW("null scene handle");
        Scene &s = *h->s;
        if (!s.committed) AM_THROW("scene not committed (call amvpt_scene_commit first)");
        if (!(o && d && maxt && prim)) AM_THROW("invalid argument: amvpt_intersect_static needs o, d, maxt and prim");
        if (n > MAX_RAYS) AM_THROW("invalid argument: amvpt_intersect_static: %u rays exceed the static scratch (AMVPT_STATIC_MAX_RAYS = %u); split the batch or use amvpt_intersect", n, MAX_RAYS);
        AM_CUDA(cudaSetDevice(s.device));
        unsigned char *scr = scratch_on(s.device);
        const size_t N = n; cudaStream_t st = nullptr;
        AM_CUDA(cudaMemsetAsync(scr + OFF_C, 0, 8 * CTR_COUNT, st));           // the only part that must start at zero
        if (n) {
            AM_CUDA(cudaMemcpyAsync(scr + OFF_O, o, N * 12, cudaMemcpyHostToDevice, st));
            AM_CUDA(cudaMemcpyAsync(scr + OFF_D, d, N * 12, cudaMemcpyHostToDevice, st));
            AM_CUDA(cudaMemcpyAsync(scr + OFF_M, maxt, N * 4, cudaMemcpyHostToDevice, st));
            intersect_static_kernel<<<(n + 127) / 128, 128, 0, st>>>(s.d, scr, n, any_hit);
            AM_CUDA(cudaGetLastError());
            AM_CUDA(cudaMemcpyAsync(prim, scr + OFF_P, N * 4, cudaMemcpyDeviceToHost, st));
            if (tuv) AM_CUDA(cudaMemcpyAsync(tuv, scr + OFF_T, N * 12, cudaMemcpyDeviceToHost, st));
        }
        unsigned long long c[CTR_COUNT] = { 0 };
        if (stats2) AM_CUDA(cudaMemcpyAsync(c, scr + OFF_C, sizeof(c), cudaMemcpyDeviceToHost, st));
        AM_CUDA(cudaStreamSynchronize(st));
        if (stats2) { stats2[0] = c[CTR_NODES]; stats2[1] = c[CTR_TRIS]; }
        return AMVPT_OK;
    });
}
