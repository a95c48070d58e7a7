// comm.cpp -- multi-GPU hand-off of the AMVPT path behind the C ABI (the reference has no distributed code at all):
// one process per GPU, NCCL over NVLink / NVSwitch for exactly two exchanges,
//   * amvpt_film_gather: every rank sends the view tiles it rendered (plus the foreign pixels just right of / below them, where the rare
//     jittered sample that rounds up into the next pixel lands) to the root, which adds them into its film -- 1 / world of the quilt per rank instead of a
//     whole-quilt reduction; with sample reuse or a reconstruction filter wider than a pixel samples land in foreign tiles, and the
//     whole quilt is sum-reduced instead;
//   * the per-view statistics all-reduce of the adaptive allocator (render.cpp), issued on the render stream.
// NCCL is bound at run time (dlopen: the copy already loaded into the process, e.g. by torch.distributed, else the system libnccl.so.2),
// so the library itself carries no link-time dependency and single-GPU users never touch it.
#include "../../include/amvpt.h"
#include "host.h"
#include <dlfcn.h>
#include <functional>
#include <set>
#include <cstring>
#include <algorithm>
#include <nccl.h>

namespace amvpt {

struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi &nccl() {
    static NcclApi api; static bool tried = false;
    if (tried) { if (!api.lib) AM_THROW("NCCL is not available: libnccl.so.2 could not be loaded"); return api; }
    tried = true;
    const char *names[] = { "libnccl.so.2", "libnccl.so" };
    for (const char *n : names) { api.lib = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL); if (api.lib) break; }      // a copy the process already holds
    for (const char *n : names) { if (api.lib) break; api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); }
    if (!api.lib) AM_THROW("NCCL is not available: %s", dlerror());
    auto sym = [&](const char *n) { void *p = dlsym(api.lib, n); if (!p) { api.lib = nullptr; AM_THROW("NCCL symbol %s not found", n); } return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId)) sym("ncclGetUniqueId"); api.CommInitRank = (decltype(api.CommInitRank)) sym("ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy)) sym("ncclCommDestroy"); api.AllReduce = (decltype(api.AllReduce)) sym("ncclAllReduce");
    api.Reduce = (decltype(api.Reduce)) sym("ncclReduce"); api.Send = (decltype(api.Send)) sym("ncclSend"); api.Recv = (decltype(api.Recv)) sym("ncclRecv");
    api.GroupStart = (decltype(api.GroupStart)) sym("ncclGroupStart"); api.GroupEnd = (decltype(api.GroupEnd)) sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString)) sym("ncclGetErrorString");
    return api;
}
#define AM_NCCL(x) do { ncclResult_t r_ = (x); if (r_ != ncclSuccess) AM_THROW("NCCL error %s: %s", #x, nccl().GetErrorString(r_)); } while (0)

struct Comm { ncclComm_t comm = nullptr; uint32_t rank = 0, world = 1; int device = 0; float4 *stage = nullptr; size_t stage_px = 0; };

void comm_allreduce_sum(void *comm, float *dev, uint32_t n, cudaStream_t st) {
    Comm *c = (Comm *) comm; if (!c || c->world <= 1) return;
    AM_NCCL(nccl().AllReduce(dev, dev, n, ncclFloat, ncclSum, c->comm, st));
}

// What one rank hands over: the view tiles it rendered, plus the film pixels just right of / below them that are not its own -- a sample
// position pixel + jitter can round up to the next pixel (float32), so a handful of samples per render land one pixel outside the tile
// (never left of or above it).  Every pixel is listed at most once; runs of such pixels become 1-pixel-wide rectangles.
static std::vector<TileDesc> gather_tiles(const Scene &s, uint32_t rank, uint32_t world, size_t *pixels) {
    Shard sh; sh.rank = rank; sh.world = world;
    std::vector<TileDesc> own = rank_tiles(s, sh, false), out = own;
    auto is_own = [&](uint32_t x, uint32_t y) { for (auto &t : own) if (x >= t.x0 && x < t.x0 + t.w && y >= t.y0 && y < t.y0 + t.h) return true; return false; };
    std::set<std::pair<uint32_t, uint32_t>> seen;
    auto run = [&](uint32_t x0, uint32_t y0, uint32_t dx, uint32_t dy, uint32_t n) {      // n pixels from (x0, y0) in direction (dx, dy); emits maximal runs of new foreign pixels
        uint32_t start = 0, len = 0;
        auto flush = [&]() { if (!len) return; TileDesc t{}; t.x0 = x0 + dx * start; t.y0 = y0 + dy * start; t.w = dx ? len : 1; t.h = dy ? len : 1; out.push_back(t); len = 0; };
        for (uint32_t i = 0; i < n; ++i) {
            const uint32_t x = x0 + dx * i, y = y0 + dy * i;
            const bool take = x < s.film_w && y < s.film_h && !is_own(x, y) && seen.insert({ x, y }).second;
            if (take) { if (!len) start = i; ++len; } else flush();
        }
        flush();
    };
    for (auto &t : own) { run(t.x0 + t.w, t.y0, 0, 1, t.h + 1); run(t.x0, t.y0 + t.h, 1, 0, t.w); }      // right column incl. the corner, bottom row
    size_t px = 0; for (auto &q : out) { q.lane_begin = px; px += (size_t) q.w * q.h; }
    *pixels = px; return out;
}

void film_gather(Scene &s, void *comm, uint32_t root, cudaStream_t st) {
    Comm *c = (Comm *) comm; if (!c || c->world <= 1) return;
    if (!s.committed) AM_THROW("film_gather: scene not committed");
    if (root >= c->world) AM_THROW("film_gather: root %u out of range", root);
    AM_CUDA(cudaSetDevice(s.device));
    const size_t npix = (size_t) s.film_w * s.film_h;
    if (s.reuse_samples() > 0 || s.rfilter_type != 0) {       // samples land in foreign tiles: sum the whole quilt
        AM_NCCL(nccl().Reduce(s.film, s.film, npix * 4, ncclFloat, ncclSum, (int) root, c->comm, st));
        return;
    }
    size_t max_px = 0; std::vector<std::vector<TileDesc>> all(c->world); std::vector<size_t> px(c->world);
    for (uint32_t r = 0; r < c->world; ++r) { all[r] = gather_tiles(s, r, c->world, &px[r]); if (r != root) max_px = std::max(max_px, px[r]); }
    const size_t need = c->rank == root ? max_px : px[c->rank];
    if (c->stage_px < need) { if (c->stage) cudaFree(c->stage); c->stage = nullptr; c->stage_px = 0; AM_CUDA(cudaMalloc((void **) &c->stage, std::max<size_t>(need, 1) * sizeof(float4))); c->stage_px = need; }
    auto tiles_to_device = [&](const std::vector<TileDesc> &t) {
        if (t.size() > s.tiles_cap) { s.tiles_cap = (uint32_t) t.size(); s.tiles_dev = (TileDesc *) s.dalloc(sizeof(TileDesc) * s.tiles_cap); }
        AM_CUDA(cudaMemcpyAsync(s.tiles_dev, t.data(), sizeof(TileDesc) * t.size(), cudaMemcpyHostToDevice, st));
    };
    if (c->rank != root) {
        if (!px[c->rank]) return;
        tiles_to_device(all[c->rank]);
        launch_film_pack(s.film, s.film_w, s.tiles_dev, (uint32_t) all[c->rank].size(), c->stage, px[c->rank], false, st);
        AM_NCCL(nccl().Send(c->stage, px[c->rank] * 4, ncclFloat, (int) root, c->comm, st));
    } else {
        for (uint32_t r = 0; r < c->world; ++r) {
            if (r == root || !px[r]) continue;
            AM_NCCL(nccl().Recv(c->stage, px[r] * 4, ncclFloat, (int) r, c->comm, st));
            // the tile list must outlive the copy: stream order keeps the previous unpack ahead of this upload
            tiles_to_device(all[r]);
            launch_film_pack(s.film, s.film_w, s.tiles_dev, (uint32_t) all[r].size(), c->stage, px[r], true, st);
        }
    }
    AM_CUDA(cudaGetLastError());
}

} // namespace amvpt

using namespace amvpt;
struct amvpt_scene { Scene *s; };
extern "C" {
extern int amvpt_guard_call(const std::function<int()> &fn);     // capi.cpp: maps exceptions to error codes + amvpt_last_error

int amvpt_comm_unique_id(uint8_t *id128) {
    return amvpt_guard_call([&] { if (!id128) AM_THROW("amvpt_comm_unique_id: null buffer"); ncclUniqueId id; AM_NCCL(nccl().GetUniqueId(&id));
        static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes"); std::memcpy(id128, &id, 128); return AMVPT_OK; });
}
int amvpt_comm_init(const uint8_t *id128, uint32_t rank, uint32_t world, int device, amvpt_comm **out) {
    return amvpt_guard_call([&] { if (!id128 || !out) AM_THROW("amvpt_comm_init: null argument"); if (!world || rank >= world) AM_THROW("amvpt_comm_init: invalid rank %u / %u", rank, world);
        AM_CUDA(cudaSetDevice(device)); ncclUniqueId id; std::memcpy(&id, id128, 128);
        Comm *c = new Comm(); c->rank = rank; c->world = world; c->device = device;
        ncclResult_t r = nccl().CommInitRank(&c->comm, (int) world, id, (int) rank);
        if (r != ncclSuccess) { delete c; AM_THROW("NCCL error ncclCommInitRank: %s", nccl().GetErrorString(r)); }
        *out = (amvpt_comm *) c; return AMVPT_OK; });
}
void amvpt_comm_destroy(amvpt_comm *h) {
    Comm *c = (Comm *) h; if (!c) return;
    cudaSetDevice(c->device); if (c->stage) cudaFree(c->stage);
    if (c->comm) { try { nccl().CommDestroy(c->comm); } catch (...) {} }
    delete c;
}
int amvpt_film_gather(amvpt_scene *s, amvpt_comm *c, uint32_t root, void *stream) {
    return amvpt_guard_call([&] { if (!s || !s->s) AM_THROW("null scene handle"); film_gather(*s->s, c, root, (cudaStream_t) stream); return AMVPT_OK; });
}
int amvpt_gather_tiles(amvpt_scene *s, uint32_t rank, uint32_t world, uint32_t *out5, uint32_t cap) {
    return amvpt_guard_call([&] { if (!s || !s->s || !out5) AM_THROW("amvpt_gather_tiles: null argument"); if (!world || rank >= world) AM_THROW("amvpt_gather_tiles: invalid rank %u / %u", rank, world);
        size_t px = 0; std::vector<TileDesc> t = gather_tiles(*s->s, rank, world, &px);
        if (t.size() > cap) AM_THROW("amvpt_gather_tiles: buffer too small (%zu tiles)", t.size());
        for (size_t i = 0; i < t.size(); ++i) { out5[5 * i] = t[i].x0; out5[5 * i + 1] = t[i].y0; out5[5 * i + 2] = t[i].w; out5[5 * i + 3] = t[i].h; out5[5 * i + 4] = (uint32_t) t[i].lane_begin; }
        return (int) t.size(); });
}
int amvpt_comm_allreduce_sum(amvpt_comm *c, float *dev, uint32_t n, void *stream) {
    return amvpt_guard_call([&] { comm_allreduce_sum(c, dev, n, (cudaStream_t) stream); return AMVPT_OK; });
}
}
