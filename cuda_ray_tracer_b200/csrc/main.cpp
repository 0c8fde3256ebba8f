// main.cpp — `raytracer <scene.txt>`: the command-line entry of the reference (main.cu:25-94) on top
// of the C ABI (include/rt_api.h).  Same positional argument, same stdout timing lines, same output
// file (the name on the scene's `png W H name` line, written into the CWD).  Everything it computes
// goes through libb200rt.so; there is no CPU rendering path.
//
// Extra flags (not in the reference):
//   --bounds shipped|computed   Morton bounds: the reference as shipped (+INF/-INF, every code 0,
//                               SURVEY F1) or the box parse.cpp accumulates and drops (default)
//   --spp ref|true              aa>=2 sample semantics; default ref
//   --near-first | --ref-order  traversal lane (RT_TRAVERSE_NEAR_FIRST / RT_TRAVERSE_REF_ORDER); default: exact near-first
//   --bits 30|63                Morton code width
//   --device N, --counters, --ids FILE (raw int32 hit ids), --repeat N
//   --gpus N                    render on N GPUs of this node (rt_comm_init + rt_scene_broadcast + rt_render_multi: scene
//                               replicated over NCCL, interleaved tiles, owned tiles gathered on GPU 0); same PNG bytes
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "rt_api.h"

static double now_s() { return std::chrono::duration<double>(std::chrono::high_resolution_clock::now().time_since_epoch()).count(); }
static void die(const char* what) {  // CUDA_CHECK → exit(EXIT_FAILURE), main.cu:14-23
  fprintf(stderr, "%s: %s\n", what, rt_last_error());
  exit(EXIT_FAILURE);
}

int main(int argc, char* argv[]) {
  if (argc < 2) { fprintf(stderr, "usage: %s scene.txt [--bounds shipped|computed] [--spp ref|true] [--near-first|--ref-order] [--bits 30|63] [--device N] [--counters] [--ids FILE] [--repeat N] [--gpus N]\n", argv[0]); return 1; }
  unsigned bflags = RT_BOUNDS_COMPUTED, rflags = RT_SPP_REF;
  int bits = 30, device = 0, repeat = 1, gpus = 1; bool counters = false; std::string ids_path;
  for (int i = 2; i < argc; i++) {
    std::string a = argv[i];
    if (a == "--bounds" && i + 1 < argc) bflags = strcmp(argv[++i], "shipped") == 0 ? RT_BOUNDS_FROM_DESC : RT_BOUNDS_COMPUTED;
    else if (a == "--spp" && i + 1 < argc) { if (strcmp(argv[++i], "true") == 0) rflags |= RT_SPP_TRUE; }
    else if (a == "--near-first") rflags |= RT_TRAVERSE_NEAR_FIRST;
    else if (a == "--ref-order") rflags |= RT_TRAVERSE_REF_ORDER;
    else if (a == "--bits" && i + 1 < argc) bits = atoi(argv[++i]);
    else if (a == "--device" && i + 1 < argc) device = atoi(argv[++i]);
    else if (a == "--counters") { counters = true; rflags |= RT_COUNTERS; }
    else if (a == "--ids" && i + 1 < argc) ids_path = argv[++i];
    else if (a == "--repeat" && i + 1 < argc) repeat = atoi(argv[++i]);
    else if (a == "--gpus" && i + 1 < argc) gpus = atoi(argv[++i]);
    else { fprintf(stderr, "unknown argument %s\n", argv[i]); return 1; }
  }
  rt_host_scene* hs = nullptr;
  if (rt_parse_scene_file(argv[1], &hs) != RT_OK) { std::cout << rt_last_error() << std::endl; exit(1); }  // parse.cpp:22-25,218-221
  rt_scene_desc d;
  rt_host_scene_desc(hs, &d);

  if (gpus > 1) {  // ---- N GPUs of this node: same stdout lines, same PNG ----
    if (!ids_path.empty()) { fprintf(stderr, "--ids is a single-GPU option\n"); return 1; }
    double t0 = now_s();
    rt_comm* comm = nullptr;
    if (rt_comm_init(gpus, nullptr, &comm) != RT_OK) die("rt_comm_init");
    std::vector<rt_scene*> scenes((size_t)gpus, nullptr);
    if (rt_scene_broadcast(comm, &d, 0, scenes.data()) != RT_OK) die("rt_scene_broadcast");
    std::cout << "Initialize raw config time: " << now_s() - t0 << " seconds" << std::endl;  // main.cu:39
    const size_t npx = (size_t)d.width * (size_t)d.height;
    std::vector<uint8_t> img(npx * 4);
    std::vector<uint8_t*> d_img((size_t)gpus, nullptr);
    t0 = now_s();
    for (int k = 0; k < gpus; k++) d_img[k] = (uint8_t*)rt_device_alloc(rt_comm_local_device(comm, k), npx * 4);
    std::cout << "Malloc and transfer to device time: " << now_s() - t0 << " seconds" << std::endl;  // main.cu:61
    rt_counters cnt; memset(&cnt, 0, sizeof(cnt));
    for (int rep = 0; rep < repeat; rep++) {
      if (d.num_prims > 0) {
        float ms = 0.f;
        if (rt_comm_build_lbvh(comm, scenes.data(), bits, bflags, &ms) != RT_OK) die("rt_build_lbvh");
        printf("LBVH Build time (N=%d): %.3f ms\n", d.num_prims, ms);
      }
      t0 = now_s();
      if (rt_render_multi(comm, scenes.data(), d.width, d.height, d.aa, rflags, 0, 0, 0, d_img.data(), img.data(), &cnt) != RT_OK) die("rt_render_multi");
      const double wall = now_s() - t0;
      std::cout << "Render time: " << cnt.render_ms * 1e-3 << " seconds" << std::endl;
      std::cout << "Transfer to host time: " << (wall - cnt.render_ms * 1e-3 > 0 ? wall - cnt.render_ms * 1e-3 : 0.0) << " seconds" << std::endl;
      if (counters)
        printf("[counters] gpus=%d rays=%llu samples=%llu nodes=%llu leaf_tests=%llu launches=%u Mrays/s=%.2f\n", gpus, (unsigned long long)cnt.rays,
               (unsigned long long)cnt.samples, (unsigned long long)cnt.nodes_visited, (unsigned long long)cnt.leaf_tests, cnt.kernel_launches,
               cnt.render_ms > 0 ? cnt.rays / (cnt.render_ms * 1e3) : 0.0);
    }
    if (rt_png_write(rt_host_scene_filename(hs), d.width, d.height, img.data()) != RT_OK) fprintf(stderr, "%s\n", rt_last_error());
    t0 = now_s();
    for (int k = 0; k < gpus; k++) { rt_device_free(rt_comm_local_device(comm, k), d_img[k]); rt_scene_destroy(scenes[k]); }
    rt_comm_destroy(comm);
    rt_host_scene_free(hs);
    std::cout << "cudaFree time: " << now_s() - t0 << " seconds" << std::endl;
    return 0;
  }

  double t0 = now_s();
  rt_scene* sc = nullptr;
  if (rt_scene_create(&d, device, &sc) != RT_OK) die("rt_scene_create");
  std::cout << "Initialize raw config time: " << now_s() - t0 << " seconds" << std::endl;  // main.cu:39

  const size_t npx = (size_t)d.width * (size_t)d.height;
  std::vector<uint8_t> img(npx * 4);
  std::vector<int32_t> ids(ids_path.empty() ? 0 : npx);
  t0 = now_s();
  if (rt_render_reserve(sc, d.width, d.height) != RT_OK) die("rt_render_reserve");  // the device image (cudaMalloc d_image + d_rawConfig, main.cu:50-60)
  std::cout << "Malloc and transfer to device time: " << now_s() - t0 << " seconds" << std::endl;  // main.cu:61
  rt_counters cnt; memset(&cnt, 0, sizeof(cnt));
  for (int rep = 0; rep < repeat; rep++) {
    if (d.num_prims > 0) {  // main.cu:44-46
      float ms = 0.f;
      if (rt_build_lbvh(sc, bits, bflags, nullptr, &ms) != RT_OK) die("rt_build_lbvh");
      printf("LBVH Build time (N=%d): %.3f ms\n", d.num_prims, ms);  // lbvh_builder.cu:489
    }
    t0 = now_s();
    if (rt_render_host(sc, d.width, d.height, d.aa, rflags, nullptr, img.data(), ids.empty() ? nullptr : ids.data(), nullptr, &cnt) != RT_OK)
      die("rt_render");
    const double wall = now_s() - t0;
    std::cout << "Render time: " << cnt.render_ms * 1e-3 << " seconds" << std::endl;  // main.cu:71 (device time of render())
    (void)wall;
    std::cout << "Transfer to host time: " << cnt.d2h_ms * 1e-3 << " seconds" << std::endl;  // main.cu:73-80: the device->host copy of the image, timed on its own
    if (counters)
      printf("[counters] rays=%llu samples=%llu nodes=%llu leaf_tests=%llu shaded=%llu iters=%u launches=%u Mrays/s=%.2f\n",
             (unsigned long long)cnt.rays, (unsigned long long)cnt.samples, (unsigned long long)cnt.nodes_visited,
             (unsigned long long)cnt.leaf_tests, (unsigned long long)cnt.shaded_hits, cnt.wavefront_iters, cnt.kernel_launches,
             cnt.render_ms > 0 ? cnt.rays / (cnt.render_ms * 1e3) : 0.0);
  }
  if (rt_png_write(rt_host_scene_filename(hs), d.width, d.height, img.data()) != RT_OK)  // main.cu:82-83
    fprintf(stderr, "%s\n", rt_last_error());  // save_image fails silently in the reference (libpng.cpp:73-107)
  if (!ids_path.empty()) { FILE* f = fopen(ids_path.c_str(), "wb"); if (f) { fwrite(ids.data(), 4, npx, f); fclose(f); } }
  t0 = now_s();
  rt_scene_destroy(sc);
  rt_host_scene_free(hs);
  std::cout << "cudaFree time: " << now_s() - t0 << " seconds" << std::endl;  // main.cu:93
  return 0;
}
