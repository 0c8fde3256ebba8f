// comm.cu — multi-GPU helpers of the C ABI: rt_comm_init / rt_comm_init_rank, rt_scene_broadcast,
// rt_comm_update_triangles, rt_framebuffer_gather, rt_render_multi.
//
// The reference is a single-GPU program (main.cu:25-94, job.slurm:7-11).  The path shards by image tiles: the scene is
// REPLICATED (NCCL broadcast of the primitive arrays from the root, once per scene; 36 bytes per triangle per frame for
// a deforming scene), every GPU builds the same LBVH itself (deterministic, cheaper than broadcasting 64 B per node),
// renders the tiles t with t % ranks == rank, and the owned tiles are gathered on the root: a pack kernel makes them
// contiguous, ncclSend / ncclRecv move 1/ranks of the image per GPU over NVLink, an unpack kernel on the root puts
// them in place.  No collective inside a frame's data path.
//
// Two ways to form the communicator: one process driving all GPUs (the C++ host: rt_comm_init -> ncclCommInitAll) or
// one process per GPU (torchrun: rt_comm_unique_id on one rank, rt_comm_init_rank everywhere).  NCCL is resolved at
// run time (dlopen libnccl.so.2: in a torch process that is torch's own copy, already loaded), so libb200rt.so has no
// NCCL dependency until a communicator is asked for.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {

struct NcclApi {
  void* so = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;

int load_nccl() {
  if (g_nccl.so) return RT_OK;
  const char* names[] = {getenv("RT_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  void* so = nullptr;
  for (const char* n : names) {
    if (!n || !*n) continue;
    so = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (so) break;
  }
  if (!so) { rt_set_error(std::string("rt_comm: cannot load NCCL (libnccl.so.2): ") + dlerror()); return RT_ERR_INVALID; }
#define RT_SYM(field, name)                                                                       \
  g_nccl.field = (decltype(g_nccl.field))dlsym(so, name);                                         \
  if (!g_nccl.field) { rt_set_error(std::string("rt_comm: NCCL symbol missing: ") + name); dlclose(so); return RT_ERR_INVALID; }
  RT_SYM(GetUniqueId, "ncclGetUniqueId") RT_SYM(CommInitAll, "ncclCommInitAll") RT_SYM(CommInitRank, "ncclCommInitRank")
  RT_SYM(CommDestroy, "ncclCommDestroy") RT_SYM(Broadcast, "ncclBroadcast") RT_SYM(Send, "ncclSend") RT_SYM(Recv, "ncclRecv")
  RT_SYM(GroupStart, "ncclGroupStart") RT_SYM(GroupEnd, "ncclGroupEnd") RT_SYM(GetErrorString, "ncclGetErrorString")
#undef RT_SYM
  g_nccl.so = so;
  return RT_OK;
}

#define RT_NCCL(call)                                                                                         \
  do {                                                                                                        \
    ncclResult_t r_ = (call);                                                                                 \
    if (r_ != ncclSuccess) {                                                                                  \
      rt_set_error(std::string(#call) + " failed: " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r_) : "?")); \
      return RT_ERR_CUDA;                                                                                     \
    }                                                                                                         \
  } while (0)

}  // namespace

struct rt_comm {
  int nranks = 0;                  // GPUs in the communicator
  std::vector<int> dev;            // local devices
  std::vector<int> rank;           // their ranks
  std::vector<ncclComm_t> comm;
  std::vector<cudaStream_t> stream;  // one internal stream per local device (scene broadcast)
  // framebuffer gather scratch, per local device: packed owned tiles; on the root: everybody else's
  std::vector<uint8_t*> pack; std::vector<size_t> pack_bytes;
  uint8_t* recv = nullptr; size_t recv_bytes = 0; int recv_local = -1;
};

namespace {

int local_of_rank(const rt_comm* c, int r) {
  for (size_t k = 0; k < c->rank.size(); k++) if (c->rank[k] == r) return (int)k;
  return -1;
}

// owned tiles of `rank` made contiguous: out[lt][y][x] (RGBA words), tile t = lt * ranks + rank.  Pixels of an edge
// tile that fall outside the image are skipped (their words are never read on the other side either).
__global__ void pack_tiles_kernel(const uint32_t* __restrict__ img, int W, int H, int tw, int th, int tiles_x, int ranks, int rank,
                                  long long nwords, uint32_t* __restrict__ out) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nwords) return;
  const int per = tw * th;
  const long long lt = g / per; const int in = (int)(g - lt * per);
  const long long t = lt * ranks + rank;
  const int tx = (int)(t % tiles_x), ty = (int)(t / tiles_x);
  const int x = tx * tw + in % tw, y = ty * th + in / tw;
  if (x < W && y < H) out[g] = img[(size_t)y * W + x];
}
__global__ void unpack_tiles_kernel(uint32_t* __restrict__ img, int W, int H, int tw, int th, int tiles_x, int ranks, int rank,
                                    long long nwords, const uint32_t* __restrict__ in_buf) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nwords) return;
  const int per = tw * th;
  const long long lt = g / per; const int in = (int)(g - lt * per);
  const long long t = lt * ranks + rank;
  const int tx = (int)(t % tiles_x), ty = (int)(t / tiles_x);
  const int x = tx * tw + in % tw, y = ty * th + in / tw;
  if (x < W && y < H) img[(size_t)y * W + x] = in_buf[g];
}

long long owned_tiles(long long ntiles, int ranks, int rank) { return ntiles > rank ? (ntiles - rank + ranks - 1) / ranks : 0; }

int finish_init(rt_comm* c) {
  c->stream.resize(c->dev.size()); c->pack.assign(c->dev.size(), nullptr); c->pack_bytes.assign(c->dev.size(), 0);
  for (size_t k = 0; k < c->dev.size(); k++) {
    RT_CUDA(cudaSetDevice(c->dev[k]));
    RT_CUDA(cudaStreamCreateWithFlags(&c->stream[k], cudaStreamNonBlocking));
  }
  return RT_OK;
}

}  // namespace

// one process, ndev GPUs (devices == NULL: 0..ndev-1); rank k = devices[k]
extern "C" int rt_comm_init(int ndev, const int* devices, rt_comm** out) {
  if (ndev < 1 || !out) { rt_set_error("rt_comm_init: bad argument"); return RT_ERR_INVALID; }
  int have = 0;
  if (cudaGetDeviceCount(&have) != cudaSuccess || have < ndev) {
    cudaGetLastError();
    rt_set_error("rt_comm_init: " + std::to_string(ndev) + " GPUs asked for, " + std::to_string(have) + " visible");
    return RT_ERR_CUDA;
  }
  int rc = load_nccl(); if (rc) return rc;
  rt_comm* c = new rt_comm();
  c->nranks = ndev;
  for (int k = 0; k < ndev; k++) { c->dev.push_back(devices ? devices[k] : k); c->rank.push_back(k); }
  c->comm.resize(ndev);
  ncclResult_t r = g_nccl.CommInitAll(c->comm.data(), ndev, c->dev.data());
  if (r != ncclSuccess) { rt_set_error(std::string("ncclCommInitAll failed: ") + g_nccl.GetErrorString(r)); delete c; return RT_ERR_CUDA; }
  rc = finish_init(c);
  if (rc) { delete c; return rc; }
  *out = c;
  return RT_OK;
}

// one process per GPU: id128 = the 128 bytes rt_comm_unique_id produced on ONE rank (the host distributes them)
extern "C" int rt_comm_unique_id(void* id128) {
  if (!id128) { rt_set_error("rt_comm_unique_id: null argument"); return RT_ERR_INVALID; }
  int rc = load_nccl(); if (rc) return rc;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  RT_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return RT_OK;
}
extern "C" int rt_comm_init_rank(int nranks, int rank, const void* id128, int device, rt_comm** out) {
  if (nranks < 1 || rank < 0 || rank >= nranks || !id128 || !out) { rt_set_error("rt_comm_init_rank: bad argument"); return RT_ERR_INVALID; }
  int rc = load_nccl(); if (rc) return rc;
  RT_CUDA(cudaSetDevice(device));
  rt_comm* c = new rt_comm();
  c->nranks = nranks; c->dev.push_back(device); c->rank.push_back(rank); c->comm.resize(1);
  ncclUniqueId id; memcpy(&id, id128, sizeof(id));
  ncclResult_t r = g_nccl.CommInitRank(&c->comm[0], nranks, id, rank);
  if (r != ncclSuccess) { rt_set_error(std::string("ncclCommInitRank failed: ") + g_nccl.GetErrorString(r)); delete c; return RT_ERR_CUDA; }
  rc = finish_init(c);
  if (rc) { delete c; return rc; }
  *out = c;
  return RT_OK;
}
extern "C" int rt_comm_destroy(rt_comm* c) {
  if (!c) return RT_OK;
  for (size_t k = 0; k < c->dev.size(); k++) {
    cudaSetDevice(c->dev[k]);
    if (k < c->stream.size() && c->stream[k]) { cudaStreamSynchronize(c->stream[k]); cudaStreamDestroy(c->stream[k]); }
    if (k < c->pack.size()) cudaFree(c->pack[k]);
    if ((int)k == c->recv_local) cudaFree(c->recv);
    if (k < c->comm.size() && c->comm[k]) g_nccl.CommDestroy(c->comm[k]);
  }
  delete c;
  return RT_OK;
}
extern "C" int rt_comm_size(const rt_comm* c) { return c ? c->nranks : 0; }
extern "C" int rt_comm_num_local(const rt_comm* c) { return c ? (int)c->dev.size() : 0; }
extern "C" int rt_comm_local_rank(const rt_comm* c, int local) { return (c && local >= 0 && local < (int)c->rank.size()) ? c->rank[local] : -1; }
extern "C" int rt_comm_local_device(const rt_comm* c, int local) { return (c && local >= 0 && local < (int)c->dev.size()) ? c->dev[local] : -1; }

// Tile shape for `ranks` GPUs.  Tile t (row-major) belongs to rank t % ranks, so when the number of tiles per image row is
// a multiple of `ranks` every rank would own whole COLUMNS of tiles and the load would follow the picture's horizontal
// structure.  Candidates are 8 rows high (two rows of the 8 x 4 pixel blocks a warp's slots run over; 1-row tiles fall
// back to row-major slots and measured ~2 % slower for the slowest rank: profiles/tile_shape_scan_r02.txt) and 32..1280
// wide; the choice minimises, in this order, the common factor of tiles-per-row and `ranks`, the padding of the last tile
// in a row, and the tile size (finer interleave = better balance): 256 x 8 at 3840 wide (15 tiles per row, diagonal
// ownership), 128 x 8 at 1920.
extern "C" void rt_tile_shape(int ranks, int width, int* tile_w, int* tile_h) {
  static const int widths[] = {32, 64, 96, 128, 160, 256, 384, 512, 640, 768, 1280};
  int bw = 32, bh = 8;
  if (ranks > 1 && width > 0) {
    long long best[3] = {1ll << 60, 0, 0};
    for (const int wdt : widths) {
      const int tx = (width + wdt - 1) / wdt;
      int a = tx, b = ranks; while (b) { const int t = a % b; a = b; b = t; }  // gcd
      const long long key[3] = {a, (long long)tx * wdt - width, wdt};
      if (std::lexicographical_compare(key, key + 3, best, best + 3)) { best[0] = key[0]; best[1] = key[1]; best[2] = key[2]; bw = wdt; }
    }
  }
  if (tile_w) *tile_w = bw;
  if (tile_h) *tile_h = bh;
}

// ---- scene replication --------------------------------------------------------------------------------------------
// root_desc: the scene on the process that owns rank `root` (host or device arrays, as for rt_scene_create); ignored
// elsewhere.  Every local device gets the same arrays over NCCL and creates its own arena from them.
extern "C" int rt_scene_broadcast(rt_comm* c, const rt_scene_desc* root_desc, int root, rt_scene** scenes_out) {
  if (!c || !scenes_out || root < 0 || root >= c->nranks) { rt_set_error("rt_scene_broadcast: bad argument"); return RT_ERR_INVALID; }
  const int nl = (int)c->dev.size(), lroot = local_of_rank(c, root);
  if (lroot >= 0 && !root_desc) { rt_set_error("rt_scene_broadcast: the root rank lives in this process and root_desc is NULL"); return RT_ERR_INVALID; }
  // 1. scalars and counts
  rt_scene_desc hdr; memset(&hdr, 0, sizeof(hdr));
  std::vector<void*> d_hdr(nl, nullptr);
  for (int k = 0; k < nl; k++) { RT_CUDA(cudaSetDevice(c->dev[k])); RT_CUDA(cudaMalloc(&d_hdr[k], sizeof(hdr))); }
  if (lroot >= 0) { RT_CUDA(cudaSetDevice(c->dev[lroot])); RT_CUDA(cudaMemcpyAsync(d_hdr[lroot], root_desc, sizeof(hdr), cudaMemcpyHostToDevice, c->stream[lroot])); }
  RT_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < nl; k++) RT_NCCL(g_nccl.Broadcast(d_hdr[k], d_hdr[k], sizeof(hdr), ncclUint8, root, c->comm[k], c->stream[k]));
  RT_NCCL(g_nccl.GroupEnd());
  RT_CUDA(cudaSetDevice(c->dev[0]));
  RT_CUDA(cudaMemcpyAsync(&hdr, d_hdr[0], sizeof(hdr), cudaMemcpyDeviceToHost, c->stream[0]));
  RT_CUDA(cudaStreamSynchronize(c->stream[0]));
  for (int k = 0; k < nl; k++) { RT_CUDA(cudaSetDevice(c->dev[k])); RT_CUDA(cudaStreamSynchronize(c->stream[k])); RT_CUDA(cudaFree(d_hdr[k])); }
  // 2. the arrays, packed back to back (256-byte aligned) in one staging buffer per device
  const size_t ns = (size_t)hdr.num_spheres, nt = (size_t)hdr.num_triangles, np = (size_t)hdr.num_prims;
  const size_t bytes[16] = {ns * 12, ns * 4, ns * 44, nt * 12, nt * 12, nt * 12, nt * 12, nt * 12, nt * 12, nt * 44, np * 8,
                            (size_t)hdr.num_suns * 24, (size_t)hdr.num_bulbs * 24, (size_t)hdr.num_planes * 84, 0, 0};
  const void* src[14] = {nullptr};
  if (lroot >= 0) {
    const void* s[14] = {root_desc->sphere_c, root_desc->sphere_r, root_desc->sphere_mat, root_desc->tri_p0, root_desc->tri_p1, root_desc->tri_p2,
                         root_desc->tri_nor, root_desc->tri_e1, root_desc->tri_e2, root_desc->tri_mat, root_desc->refs, root_desc->suns, root_desc->bulbs,
                         root_desc->planes};
    memcpy(src, s, sizeof(s));
  }
  size_t off[15]; off[0] = 0;
  for (int a = 0; a < 14; a++) off[a + 1] = off[a] + ((bytes[a] + 255) & ~(size_t)255);
  const size_t total = std::max<size_t>(off[14], 256);
  std::vector<char*> stage(nl, nullptr);
  for (int k = 0; k < nl; k++) { RT_CUDA(cudaSetDevice(c->dev[k])); RT_CUDA(cudaMalloc((void**)&stage[k], total)); }
  if (lroot >= 0) {
    RT_CUDA(cudaSetDevice(c->dev[lroot]));
    for (int a = 0; a < 14; a++)
      if (bytes[a]) {
        if (!src[a]) { rt_set_error("rt_scene_broadcast: a scene array is NULL while its count is not 0"); return RT_ERR_INVALID; }
        RT_CUDA(cudaMemcpyAsync(stage[lroot] + off[a], src[a], bytes[a], cudaMemcpyDefault, c->stream[lroot]));
      }
  }
  RT_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < nl; k++) RT_NCCL(g_nccl.Broadcast(stage[k], stage[k], total, ncclUint8, root, c->comm[k], c->stream[k]));
  RT_NCCL(g_nccl.GroupEnd());
  for (int k = 0; k < nl; k++) { RT_CUDA(cudaSetDevice(c->dev[k])); RT_CUDA(cudaStreamSynchronize(c->stream[k])); }
  // 3. one arena per device from the device-resident arrays
  int rc = RT_OK;
  for (int k = 0; k < nl && rc == RT_OK; k++) {
    rt_scene_desc d = hdr;
    char* b = stage[k];
    auto at = [&](int a) -> const float* { return bytes[a] ? (const float*)(b + off[a]) : nullptr; };
    d.sphere_c = at(0); d.sphere_r = at(1); d.sphere_mat = at(2); d.tri_p0 = at(3); d.tri_p1 = at(4); d.tri_p2 = at(5);
    d.tri_nor = at(6); d.tri_e1 = at(7); d.tri_e2 = at(8); d.tri_mat = at(9); d.refs = (const int32_t*)at(10);
    d.suns = at(11); d.bulbs = at(12); d.planes = at(13);
    scenes_out[k] = nullptr;
    rc = rt_scene_create(&d, c->dev[k], &scenes_out[k]);
  }
  const std::string keep = rc ? rt_last_error() : "";
  for (int k = 0; k < nl; k++) { cudaSetDevice(c->dev[k]); cudaFree(stage[k]); }
  if (rc) { for (int k = 0; k < nl; k++) if (scenes_out[k]) { rt_scene_destroy(scenes_out[k]); scenes_out[k] = nullptr; } rt_set_error(keep); }
  return rc;
}

// New vertex positions of a deforming scene: 36 bytes per triangle from the root to every GPU, then the Triangle
// constructor on each device (rt_scene_update_triangles).  p0/p1/p2: host or device pointers on the root's process
// (ignored elsewhere).  streams: one cudaStream_t per local device, or NULL for the default streams.  Asynchronous.
extern "C" int rt_comm_update_triangles(rt_comm* c, rt_scene* const* scenes, const float* p0, const float* p1, const float* p2, int root,
                                        void* const* streams) {
  if (!c || !scenes || root < 0 || root >= c->nranks) { rt_set_error("rt_comm_update_triangles: bad argument"); return RT_ERR_INVALID; }
  const int nl = (int)c->dev.size(), lroot = local_of_rank(c, root);
  if (lroot >= 0 && (!p0 || !p1 || !p2)) { rt_set_error("rt_comm_update_triangles: null vertex array on the root"); return RT_ERR_INVALID; }
  const float* src[3] = {p0, p1, p2};
  for (int k = 0; k < nl; k++) {
    rt_scene* s = scenes[k];
    if (!s) { rt_set_error("rt_comm_update_triangles: null scene"); return RT_ERR_INVALID; }
    if (s->nt == 0) return RT_OK;
    RT_CUDA(cudaSetDevice(c->dev[k]));
    cudaStream_t st = streams ? (cudaStream_t)streams[k] : nullptr;
    rt_arm_fork(s, st);  // RT_RENDER_OVERLAP_BUILD: the frame starts here, before the vertices arrive
    for (int a = 0; a < 3; a++) {
      if (!s->upd_stage[a]) RT_CUDA(cudaMalloc((void**)&s->upd_stage[a], (size_t)s->nt * 12));
      if (k == lroot && src[a] != s->upd_stage[a]) RT_CUDA(cudaMemcpyAsync(s->upd_stage[a], src[a], (size_t)s->nt * 12, cudaMemcpyDefault, st));
    }
  }
  RT_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < nl; k++)
    for (int a = 0; a < 3; a++)
      RT_NCCL(g_nccl.Broadcast(scenes[k]->upd_stage[a], scenes[k]->upd_stage[a], (size_t)scenes[k]->nt * 12, ncclUint8, root, c->comm[k],
                               streams ? (cudaStream_t)streams[k] : nullptr));
  RT_NCCL(g_nccl.GroupEnd());
  for (int k = 0; k < nl; k++) {
    rt_scene* s = scenes[k];
    const int rc = rt_scene_update_triangles(s, s->upd_stage[0], s->upd_stage[1], s->upd_stage[2], streams ? streams[k] : nullptr);
    if (rc) return rc;
  }
  return RT_OK;
}

// ---- framebuffer gather ---------------------------------------------------------------------------------------------
// d_rgba[k]: the full-size RGBA8 image of local device k holding the tiles its rank rendered (rt_render with the tile
// set {tile_w, tile_h, ranks, rank}).  After the call the image of rank `root` holds every tile.  Each non-root GPU
// sends only its own tiles (1/ranks of the image), packed; asynchronous on `streams` (NULL: the default streams).
extern "C" int rt_framebuffer_gather(rt_comm* c, uint8_t* const* d_rgba, int w, int h, int tile_w, int tile_h, int root, void* const* streams) {
  if (!c || !d_rgba || w <= 0 || h <= 0 || tile_w <= 0 || tile_h <= 0 || root < 0 || root >= c->nranks) {
    rt_set_error("rt_framebuffer_gather: bad argument"); return RT_ERR_INVALID;
  }
  if (c->nranks == 1) return RT_OK;
  const int nl = (int)c->dev.size(), R = c->nranks, lroot = local_of_rank(c, root);
  const int tw = std::min(tile_w, w), th = std::min(tile_h, h);   // as rt_render clamps them
  const int tiles_x = (w + tw - 1) / tw, tiles_y = (h + th - 1) / th;
  const long long ntiles = (long long)tiles_x * tiles_y, per = (long long)tw * th;
  auto words_of = [&](int r) { return owned_tiles(ntiles, R, r) * per; };
  std::vector<size_t> roff((size_t)R + 1, 0);
  for (int r = 0; r < R; r++) roff[r + 1] = roff[r] + (r == root ? 0 : (size_t)words_of(r) * 4);
  for (int k = 0; k < nl; k++) {
    RT_CUDA(cudaSetDevice(c->dev[k]));
    cudaStream_t st = streams ? (cudaStream_t)streams[k] : nullptr;
    if (k == lroot) {
      if (roff[R] > c->recv_bytes || c->recv_local != k) {
        if (c->recv) cudaFree(c->recv);
        c->recv = nullptr; c->recv_bytes = 0;
        RT_CUDA(cudaMalloc((void**)&c->recv, std::max<size_t>(roff[R], 256)));
        c->recv_bytes = std::max<size_t>(roff[R], 256); c->recv_local = k;
      }
      continue;
    }
    const long long nw = words_of(c->rank[k]);
    if ((size_t)nw * 4 > c->pack_bytes[k]) {
      if (c->pack[k]) cudaFree(c->pack[k]);
      c->pack[k] = nullptr; c->pack_bytes[k] = 0;
      RT_CUDA(cudaMalloc((void**)&c->pack[k], std::max<size_t>((size_t)nw * 4, 256)));
      c->pack_bytes[k] = std::max<size_t>((size_t)nw * 4, 256);
    }
    if (nw > 0) pack_tiles_kernel<<<(unsigned)((nw + 255) / 256), 256, 0, st>>>((const uint32_t*)d_rgba[k], w, h, tw, th, tiles_x, R, c->rank[k], nw, (uint32_t*)c->pack[k]);
    RT_CUDA(cudaGetLastError());
  }
  RT_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < nl; k++) {
    cudaStream_t st = streams ? (cudaStream_t)streams[k] : nullptr;
    if (k == lroot) {
      for (int r = 0; r < R; r++)
        if (r != root && words_of(r) > 0) RT_NCCL(g_nccl.Recv(c->recv + roff[r], (size_t)words_of(r) * 4, ncclUint8, r, c->comm[k], st));
    } else if (words_of(c->rank[k]) > 0) RT_NCCL(g_nccl.Send(c->pack[k], (size_t)words_of(c->rank[k]) * 4, ncclUint8, root, c->comm[k], st));
  }
  RT_NCCL(g_nccl.GroupEnd());
  if (lroot >= 0) {
    RT_CUDA(cudaSetDevice(c->dev[lroot]));
    cudaStream_t st = streams ? (cudaStream_t)streams[lroot] : nullptr;
    for (int r = 0; r < R; r++) {
      const long long nw = words_of(r);
      if (r == root || nw == 0) continue;
      unpack_tiles_kernel<<<(unsigned)((nw + 255) / 256), 256, 0, st>>>((uint32_t*)d_rgba[lroot], w, h, tw, th, tiles_x, R, r, nw, (const uint32_t*)(c->recv + roff[r]));
    }
    RT_CUDA(cudaGetLastError());
  }
  return RT_OK;
}

// ---- one frame on every local GPU -------------------------------------------------------------------------------------
// rt_build_lbvh on every local scene (each GPU builds the same tree itself), one host thread per GPU
extern "C" int rt_comm_build_lbvh(rt_comm* c, rt_scene* const* scenes, int morton_bits, unsigned flags, float* ms_out) {
  if (!c || !scenes) { rt_set_error("rt_comm_build_lbvh: bad argument"); return RT_ERR_INVALID; }
  const int nl = (int)c->dev.size();
  std::vector<int> rc(nl, RT_OK); std::vector<std::string> err(nl); std::vector<float> ms(nl, 0.f);
  std::vector<std::thread> th;
  for (int k = 0; k < nl; k++)
    th.emplace_back([&, k]() { rc[k] = rt_build_lbvh(scenes[k], morton_bits, flags, nullptr, &ms[k]); if (rc[k]) err[k] = rt_last_error(); });
  for (auto& t : th) t.join();
  for (int k = 0; k < nl; k++) if (rc[k]) { rt_set_error(err[k]); return rc[k]; }
  if (ms_out) *ms_out = *std::max_element(ms.begin(), ms.end());
  return RT_OK;
}

// render(): every local GPU renders its tiles (one host thread per GPU), the tiles are gathered on rank `root`, and the
// root's image is copied to h_rgba when that pointer is given (in the process that owns the root).  d_rgba[k]: device
// image of local GPU k.  tile_w/tile_h <= 0: rt_tile_shape.  counters: summed over the local GPUs (render_ms: the slowest).
extern "C" int rt_render_multi(rt_comm* c, rt_scene* const* scenes, int w, int h, int aa, unsigned flags, int tile_w, int tile_h, int root,
                               uint8_t* const* d_rgba, uint8_t* h_rgba, rt_counters* counters) {
  if (!c || !scenes || !d_rgba || root < 0 || root >= c->nranks) { rt_set_error("rt_render_multi: bad argument"); return RT_ERR_INVALID; }
  const int nl = (int)c->dev.size();
  if (tile_w <= 0 || tile_h <= 0) rt_tile_shape(c->nranks, w, &tile_w, &tile_h);
  std::vector<int> rc(nl, RT_OK); std::vector<std::string> err(nl); std::vector<rt_counters> cnt(nl);
  std::vector<std::thread> th;
  for (int k = 0; k < nl; k++)
    th.emplace_back([&, k]() {
      rt_tile_set ts = {tile_w, tile_h, c->nranks, c->rank[k]};
      rc[k] = rt_render(scenes[k], w, h, aa, flags, c->nranks > 1 ? &ts : nullptr, d_rgba[k], nullptr, nullptr, &cnt[k], nullptr);
      if (rc[k]) err[k] = rt_last_error();
    });
  for (auto& t : th) t.join();
  for (int k = 0; k < nl; k++) if (rc[k]) { rt_set_error(err[k]); return rc[k]; }
  int g = rt_framebuffer_gather(c, d_rgba, w, h, tile_w, tile_h, root, nullptr);
  if (g) return g;
  const int lroot = local_of_rank(c, root);
  if (lroot >= 0) {
    RT_CUDA(cudaSetDevice(c->dev[lroot]));
    if (h_rgba) RT_CUDA(cudaMemcpy(h_rgba, d_rgba[lroot], (size_t)w * h * 4, cudaMemcpyDeviceToHost));  // main.cu:77
    else RT_CUDA(cudaStreamSynchronize(nullptr));
  }
  if (counters) {
    memset(counters, 0, sizeof(*counters));
    for (int k = 0; k < nl; k++) {
      counters->rays += cnt[k].rays; counters->samples += cnt[k].samples; counters->nodes_visited += cnt[k].nodes_visited;
      counters->leaf_tests += cnt[k].leaf_tests; counters->shaded_hits += cnt[k].shaded_hits; counters->fallback_rays += cnt[k].fallback_rays;
      counters->packet_steps += cnt[k].packet_steps; counters->kernel_launches += cnt[k].kernel_launches; counters->traverse_launches += cnt[k].traverse_launches;
      counters->wavefront_iters = std::max(counters->wavefront_iters, cnt[k].wavefront_iters); counters->passes = std::max(counters->passes, cnt[k].passes);
      counters->render_ms = std::max(counters->render_ms, cnt[k].render_ms); counters->traverse_ms = std::max(counters->traverse_ms, cnt[k].traverse_ms);
    }
  }
  return RT_OK;
}

// device memory for a host program that does not link the CUDA runtime itself (the `raytracer` CLI's per-GPU images)
extern "C" void* rt_device_alloc(int device, size_t bytes) {
  void* p = nullptr;
  if (cudaSetDevice(device) != cudaSuccess || cudaMalloc(&p, bytes ? bytes : 1) != cudaSuccess) {
    rt_set_error(std::string("rt_device_alloc: ") + cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  return p;
}
extern "C" void rt_device_free(int device, void* p) { if (p && cudaSetDevice(device) == cudaSuccess) cudaFree(p); }
