// fs2_render — headless replacement of the reference's main.cpp for the path-tracing hot
// path (reference source/main.cpp:47-168 opens a Win32/OpenGL window, which is unavailable
// on this box).  Same options (-f file, -w width, -h height; first-letter matching like
// source/main.cpp:61-76) plus --spp, --seed, --device, --out, --focus x,y (the click-to-focus
// of source/main.cpp:144-153 through TraceRange), --fast (Camera::fast_render_enabled) and
// --checkpoint / --resume (the accumulation film on disk: a render continues where it stopped).
// --post-out image.ppm writes the film after the device post-processing chain (--denoise passes,
// --bloom strength[,threshold[,radius]], --flare strength[,ghosts]).
// Written against the mirrored API of host/fs2_host.h only.
//
//   fs2_render -f scene.scene -w 1920 -h 1080 --spp 64 --out image.ppm [--pfm mean.pfm] [--exr mean.exr]
#include <chrono>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "fs2_host.h"

using namespace base;

static bool write_ppm(const std::string& path, DisplayFrame* f) {
  FILE* fp = fopen(path.c_str(), "wb");
  if (!fp) return false;
  uint32 w = f->GetWidth(), h = f->GetHeight();
  fprintf(fp, "P6\n%u %u\n255\n", w, h);
  const uint8* rgb = f->GetDisplayBuffer();
  for (uint32 y = 0; y < h; y++)  // row 0 of the display buffer is the bottom scanline
    fwrite(rgb + 3 * (size_t)(h - 1 - y) * w, 1, 3 * (size_t)w, fp);
  fclose(fp);
  return true;
}

static bool write_ppm_bytes(const std::string& path, const uint8* rgb, uint32 w, uint32 h) {
  FILE* fp = fopen(path.c_str(), "wb");
  if (!fp) return false;
  fprintf(fp, "P6\n%u %u\n255\n", w, h);
  for (uint32 y = 0; y < h; y++) fwrite(rgb + 3 * (size_t)(h - 1 - y) * w, 1, 3 * (size_t)w, fp);
  return fclose(fp) == 0;
}

static bool write_pfm(const std::string& path, DisplayFrame* f) {
  FILE* fp = fopen(path.c_str(), "wb");
  if (!fp) return false;
  uint32 w = f->GetWidth(), h = f->GetHeight();
  fprintf(fp, "PF\n%u %u\n-1.0\n", w, h);  // little endian, bottom-up rows like the film
  fwrite(f->GetRenderTarget(), 4, 3 * (size_t)w * h, fp);
  fclose(fp);
  return true;
}

int main(int argc, char** argv) {
  std::string file, out = "fs2_render.ppm", pfm, exr, checkpoint, resume;
  uint32 width = 800, height = 480, spp = 16;  // 800x480: source/main.cpp:53-54
  uint64 seed = 1;
  int device = 0;
  std::vector<int> devices;  // --gpus N / --devices a,b,...: multi-device context
  int shard_mode = 0;        // 0 automatic, 1 samples, 2 tiles
  bool fast = false, have_focus = false;
  PostProcess post;
  std::string post_out;
  float fx = 0, fy = 0;
  for (int i = 1; i < argc; i++) {
    const char* a = argv[i];
    auto next = [&]() -> const char* { return i + 1 < argc ? argv[++i] : ""; };
    if (!strcmp(a, "--spp")) spp = (uint32)atoi(next());
    else if (!strcmp(a, "--seed")) seed = strtoull(next(), nullptr, 10);
    else if (!strcmp(a, "--device")) device = atoi(next());
    else if (!strcmp(a, "--gpus")) {  // the first N devices of the box
      devices.clear();
      for (int d = 0, n = atoi(next()); d < n; d++) devices.push_back(d);
    } else if (!strcmp(a, "--devices")) {  // explicit list: 0,2,5
      devices.clear();
      for (const char* p = next(); *p;) {
        devices.push_back((int)strtol(p, (char**)&p, 10));
        if (*p == ',') p++;
      }
    } else if (!strcmp(a, "--shard")) {
      const char* m = next();
      shard_mode = !strcmp(m, "samples") ? 1 : !strcmp(m, "tiles") ? 2 : 0;
    }
    else if (!strcmp(a, "--out")) out = next();
    else if (!strcmp(a, "--pfm")) pfm = next();
    else if (!strcmp(a, "--exr")) exr = next();
    else if (!strcmp(a, "--fast")) fast = true;
    else if (!strcmp(a, "--checkpoint")) checkpoint = next();
    else if (!strcmp(a, "--resume")) resume = next();
    else if (!strcmp(a, "--post-out")) post_out = next();
    else if (!strcmp(a, "--denoise")) post.denoise_passes = atoi(next());
    else if (!strcmp(a, "--bloom"))  // strength[,threshold[,radius]]
      sscanf(next(), "%f,%f,%d", &post.bloom_strength, &post.bloom_threshold, &post.bloom_radius);
    else if (!strcmp(a, "--flare")) {  // strength[,ghosts]
      post.flare_ghosts = 3;
      sscanf(next(), "%f,%d", &post.flare_strength, &post.flare_ghosts);
    }
    else if (!strcmp(a, "--focus")) have_focus = sscanf(next(), "%f,%f", &fx, &fy) == 2;
    else if (a[0] == '-' && (a[1] == 'f' || (a[1] == '-' && a[2] == 'f'))) file = next();
    else if (a[0] == '-' && (a[1] == 'w' || (a[1] == '-' && a[2] == 'w'))) width = (uint32)atoi(next());
    else if (a[0] == '-' && (a[1] == 'h' || (a[1] == '-' && a[2] == 'h'))) height = (uint32)atoi(next());
    else {
      fprintf(stderr, "unknown option %s\n", a);
      return 2;
    }
  }
  if (file.empty()) {  // source/main.cpp:78-81
    fprintf(stderr, "usage: fs2_render -f <scene file> [-w W] [-h H] [--spp N] [--seed S] "
                    "[--device D | --gpus N | --devices a,b,..] [--shard auto|samples|tiles] [--out image.ppm] [--pfm mean.pfm] [--exr mean.exr] [--focus x,y] [--fast] "
                    "[--resume film.ckpt] [--checkpoint film.ckpt] [--post-out image.ppm "
                    "[--denoise passes] [--bloom strength[,threshold[,radius]]] "
                    "[--flare strength[,ghosts]]]\n");
    return 2;
  }
  Scene scene;
  if (!scene.LoadScene(file)) return 1;
  InitializeMaterials();
  scene.Optimize();
  // default viewpoint of source/main.cpp:87-88 when the file has no camera block
  Camera cam(vector3(-5.80f, 7.05f, -47.06f), vector3(0.0f, 8.94f, 0.0f));
  if (scene.GetCameraCount()) cam = *scene.GetCamera(0);
  cam.fast_render_enabled = fast;
  DisplayFrame frame(width, height);
  if (devices.size() > 1) {
    SetTraceDevices(devices.data(), (int)devices.size());
    SetTraceSharding(shard_mode);
  } else {
    SetTraceDevice(devices.empty() ? device : devices[0]);
  }
  SetTraceSeed(seed);
  if (have_focus) {  // source/main.cpp:144-153
    cam.focal_depth = TraceRange(cam, &scene, &frame, fx, fy);
    printf("focal depth at (%.1f, %.1f): %.4f\n", fx, fy, cam.focal_depth);
  }
  if (!resume.empty() && !LoadFilmCheckpoint(&frame, resume.c_str())) {
    fprintf(stderr, "cannot resume from %s: %s\n", resume.c_str(), LastTraceError());
    return 1;
  }
  auto t0 = std::chrono::steady_clock::now();
  if (TraceSamples(cam, &scene, &frame, spp) != 0) {
    fprintf(stderr, "render failed: %s\n", LastTraceError());
    return 1;
  }
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  printf("%ux%u, %u spp on %d GPU%s in %.3f s (%.2f Mpaths/s, scene upload included)\n", width, height,
         spp, GetTraceDeviceCount(), GetTraceDeviceCount() > 1 ? "s" : "", sec,
         (double)width * height * spp / sec / 1e6);
  if (!checkpoint.empty() && !SaveFilmCheckpoint(&frame, checkpoint.c_str())) {
    fprintf(stderr, "cannot write %s: %s\n", checkpoint.c_str(), LastTraceError());
    return 1;
  }
  if (!write_ppm(out, &frame)) {
    fprintf(stderr, "cannot write %s\n", out.c_str());
    return 1;
  }
  if (!post_out.empty()) {  // denoise / bloom / flare of the finished film, on the device
    std::vector<uint8> bytes(3 * (size_t)width * height);
    if (!PostProcessFrame(&frame, post, nullptr, bytes.data())) {
      fprintf(stderr, "post-processing failed: %s\n", LastTraceError());
      return 1;
    }
    if (!write_ppm_bytes(post_out, bytes.data(), width, height)) {
      fprintf(stderr, "cannot write %s\n", post_out.c_str());
      return 1;
    }
  }
  if (!exr.empty() && !SaveFrameEXR(&frame, exr.c_str())) {
    fprintf(stderr, "cannot write %s\n", exr.c_str());
    return 1;
  }
  if (!pfm.empty() && !write_pfm(pfm, &frame)) {
    fprintf(stderr, "cannot write %s\n", pfm.c_str());
    return 1;
  }
  return 0;
}
