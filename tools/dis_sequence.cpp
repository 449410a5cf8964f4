// tools/dis_sequence.cpp — headless sequence driver with the reference's command line.
//
// Mirrors the argument handling and file naming of /root/reference/src/main.cpp:60-111 and the outputs
// of :198-203, on top of the C ABI (host C++ only; link libdis_b200.so):
//
//   dis_sequence                                   defaults: folder alley_1, frames 1..50
//   dis_sequence folder start_num_image end_num_image
//   dis_sequence folder start end max_iter patch_size coarsest_scale finest_scale patch_overlap patch_norm draw_grid
//
// Differences: frames are binary PGM (`folder/frame_%04d.pgm`, P5, maxval 255) because no PNG codec is
// linked here (convert once with any image tool); the colour-coded flow is written as PPM
// (`OF_folder/frame_%04d.ppm`) next to a Middlebury `.flo`; no windows (imshow/waitKey/draw_grid).  All
// pairs of the sequence run as ONE device call (dis_run_sequence_u8): every frame's pyramid is built once.
#include <sys/stat.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "dis_b200.h"

static bool read_pgm(const std::string& path, std::vector<uint8_t>& pix, int& w, int& h) {
  FILE* f = std::fopen(path.c_str(), "rb");
  if (!f) return false;
  char magic[3] = {0, 0, 0};
  int maxval = 0, got = 0, vals[3];
  if (std::fscanf(f, "%2s", magic) != 1 || magic[0] != 'P' || magic[1] != '5') { std::fclose(f); return false; }
  while (got < 3) {                       // width, height, maxval with optional # comments
    int c = std::fgetc(f);
    if (c == '#') { while (c != '\n' && c != EOF) c = std::fgetc(f); continue; }
    if (c == EOF) { std::fclose(f); return false; }
    if (c == ' ' || c == '\t' || c == '\n' || c == '\r') continue;
    std::ungetc(c, f);
    if (std::fscanf(f, "%d", &vals[got]) != 1) { std::fclose(f); return false; }
    ++got;
  }
  std::fgetc(f);                          // the single whitespace after maxval
  w = vals[0]; h = vals[1]; maxval = vals[2];
  if (w <= 0 || h <= 0 || maxval != 255) { std::fclose(f); return false; }
  pix.resize((size_t)w * h);
  const bool ok = std::fread(pix.data(), 1, pix.size(), f) == pix.size();
  std::fclose(f);
  return ok;
}

static std::string frame_name(const std::string& folder, int i, const char* ext) {
  char buf[32];
  std::snprintf(buf, sizeof buf, "/frame_%04d.%s", i, ext);   // main.cpp:108-109
  return folder + buf;
}

int main(int argc, char** argv) {
  std::string folder = "alley_1";          // main.cpp:63-72
  int start_num_img = 1, end_num_img = 50;
  dis_params p;
  dis_default_params(&p);
  if (argc == 1) {
  } else if (argc == 4 || argc == 11) {    // main.cpp:76-92
    folder = argv[1]; start_num_img = std::atoi(argv[2]); end_num_img = std::atoi(argv[3]);
    if (argc == 11) {
      p.max_iter = std::atoi(argv[4]); p.patch_size = std::atoi(argv[5]);
      p.coarsest_scale = std::atoi(argv[6]); p.finest_scale = std::atoi(argv[7]);
      p.patch_overlap = (float)std::atof(argv[8]); p.patch_norm = std::atoi(argv[9]);
    }
  } else {
    std::printf("Not good parameters!\n"
                "1. Run without parameters with default parameters (alley_1) from 1 to 50\n"
                "2. Set folder and images with default parameters:\n"
                "dis_sequence folder start_num_image end_num_image\n"
                "3. Add full settings\n"
                "dis_sequence folder start_num_image end_num_image max_iter patch_size coarsest_scale finest_scale "
                "patch_overlap patch_norm draw_grid\n");
    return 0;
  }
  const int n_frames = end_num_img - start_num_img + 1;
  if (n_frames < 2) return 0;              // the reference's loop body never runs
  int w = 0, h = 0;
  std::vector<uint8_t> frames, one;
  for (int i = 0; i < n_frames; ++i) {
    int wi, hi;
    if (!read_pgm(frame_name(folder, start_num_img + i, "pgm"), one, wi, hi) || (i && (wi != w || hi != h))) {
      std::printf("No image data \n");     // main.cpp:118-122
      return -1;
    }
    if (!i) { w = wi; h = hi; frames.resize((size_t)w * h * n_frames); }
    std::copy(one.begin(), one.end(), frames.begin() + (size_t)i * w * h);
  }
  dis_engine* e = nullptr;
  const char* dev = std::getenv("DIS_DEVICE");
  if (dis_create(&p, dev ? std::atoi(dev) : 0, &e) != DIS_OK) { std::fprintf(stderr, "%s\n", dis_last_error()); return 1; }
  std::vector<float> flows((size_t)2 * w * h * (n_frames - 1));
  if (dis_run_sequence_u8(e, frames.data(), n_frames, w, h, w, (size_t)w * h, flows.data()) != DIS_OK) {
    std::fprintf(stderr, "%s\n", dis_last_error());
    dis_destroy(e);
    return 1;
  }
  const std::string out_folder = "OF_" + folder;   // main.cpp:104-105
  mkdir(out_folder.c_str(), 0777);
  std::vector<uint8_t> bgr((size_t)3 * w * h), rgb(bgr.size());
  for (int i = 0; i < n_frames - 1; ++i) {
    std::printf("start %s\n", frame_name(folder, start_num_img + i, "pgm").c_str());
    const float* flow = flows.data() + (size_t)2 * w * h * i;
    if (dis_write_flo(frame_name(out_folder, start_num_img + i, "flo").c_str(), flow, w, h) != DIS_OK ||
        dis_color_code(e, flow, w, h, -1.f, bgr.data()) != DIS_OK) {   // draw_optical_flow(flowout, dst2), main.cpp:200-201
      // (from the host flow: the engine keeps only the last chunk of <= 128 pairs of a long sequence on the device)
      std::fprintf(stderr, "%s\n", dis_last_error());
      dis_destroy(e);
      return 1;
    }
    for (size_t k = 0; k < bgr.size(); k += 3) { rgb[k] = bgr[k + 2]; rgb[k + 1] = bgr[k + 1]; rgb[k + 2] = bgr[k]; }
    FILE* f = std::fopen(frame_name(out_folder, start_num_img + i, "ppm").c_str(), "wb");
    if (!f) { dis_destroy(e); return 1; }
    std::fprintf(f, "P6\n%d %d\n255\n", w, h);
    std::fwrite(rgb.data(), 1, rgb.size(), f);
    std::fclose(f);
    std::printf("finish %s\n", frame_name(folder, start_num_img + i, "pgm").c_str());
  }
  dis_destroy(e);
  return 0;
}
