// tests/cpp/mirror_test.cpp — exercises include/dis_b200.hpp exactly the way the reference's own
// code uses its classes (src/main.cpp:184-189 for OpticalFlowClass; src/optical_flow.cpp:61-90 for
// the PatchGrid level loop; src/patch_grid.cpp:68-106 for Patch).  Host C++ only; driven by
// tests/test_cpp_mirror.py, which prepares the pyramids with the oracle and compares the outputs.
//
//   mirror_test in.bin out.bin
// in.bin : int32 W,H,pad,coarsest,finest,iters,ps,patnorm; float overlap; 6 pyramids x (coarsest+1) padded planes
// out.bin: flow by constructor (2*w*h of the finest level) | flow by PatchGrid loop | K, then K x (id, ux, uy) by Patch
#include <cstdio>
#include <vector>

#include "dis_b200.hpp"

using namespace OpticalFlow;

int main(int argc, char** argv) {
  if (argc < 3) { std::fprintf(stderr, "usage: mirror_test in.bin out.bin\n"); return 2; }
  FILE* f = std::fopen(argv[1], "rb");
  if (!f) return 2;
  int hdr[8]; float overlap;
  if (std::fread(hdr, 4, 8, f) != 8 || std::fread(&overlap, 4, 1, f) != 1) return 2;
  const int W = hdr[0], H = hdr[1], pad = hdr[2], coarsest = hdr[3], finest = hdr[4], iters = hdr[5], ps = hdr[6];
  const bool patnorm = hdr[7] != 0;
  std::vector<std::vector<std::vector<float>>> pyr(6, std::vector<std::vector<float>>(coarsest + 1));
  std::vector<std::vector<float*>> ptr(6, std::vector<float*>(coarsest + 1));
  for (int k = 0; k < 6; ++k)
    for (int s = 0; s <= coarsest; ++s) {
      const size_t n = (size_t)((W >> s) + 2 * pad) * ((H >> s) + 2 * pad);
      pyr[k][s].resize(n);
      if (std::fread(pyr[k][s].data(), 4, n, f) != n) return 2;
      ptr[k][s] = pyr[k][s].data();
    }
  std::fclose(f);
  const int wf = W >> finest, hf = H >> finest;
  try {
    // (1) the constructor contract, as src/main.cpp:184-189 calls it
    std::vector<float> flow_ctor(2 * (size_t)wf * hf);
    OpticalFlowClass ofc(ptr[0].data(), ptr[1].data(), ptr[2].data(), ptr[3].data(), ptr[4].data(), ptr[5].data(), pad,
                         flow_ctor.data(), W, H, coarsest, finest, iters, ps, overlap, patnorm, false);

    // (2) the same run spelled out with PatchGrid, level by level
    fix_parameters fp;
    fp.coarsest_scale = coarsest; fp.finest_scale = finest; fp.patch_size = ps; fp.iterations = iters;
    fp.patch_normalization = patnorm;
    fp.outlierthresh = (float)ps / 2;
    fp.steps = (int)std::floor((float)ps * (1 - overlap)); if (fp.steps < 1) fp.steps = 1;
    fp.num_points_patch = ps * ps;
    std::vector<std::vector<float>> flows(coarsest + 1);
    std::vector<image_parameters> ips(coarsest + 1);
    std::vector<float> patch_u;
    std::vector<int> patch_id;
    for (int scale = coarsest; scale >= finest; --scale) {
      image_parameters& ip = ips[scale];
      ip.width = W >> scale; ip.height = H >> scale; ip.img_padding = pad;
      ip.tmp_lb = -(float)ps / 2;
      ip.tmp_ub_w = (float)(ip.width + ps / 2 - 2); ip.tmp_ub_h = (float)(ip.height + ps / 2 - 2);
      ip.tmp_w = ip.width + 2 * pad; ip.tmp_h = ip.height + 2 * pad;
      flows[scale].resize(2 * (size_t)ip.width * ip.height);
      PatchGrid grid(&ip, &fp);
      grid.init_grid(ptr[0][scale], ptr[1][scale], ptr[2][scale]);
      grid.set_target_image(ptr[3][scale], ptr[4][scale], ptr[5][scale]);
      if (scale < coarsest) grid.patch_init_from_prev_flow(flows[scale + 1].data());
      grid.inverse_search();
      grid.densification_and_create_dance_flow(flows[scale].data());

      // (3) a few patches of the coarsest level one by one through Patch
      if (scale == coarsest) {
        const int n = grid.get_num_all_patch();
        PlaneMap m1(ptr[0][scale], ip.tmp_h, ip.tmp_w), mdx(ptr[1][scale], ip.tmp_h, ip.tmp_w), mdy(ptr[2][scale], ip.tmp_h, ip.tmp_w);
        PlaneMap m2(ptr[3][scale], ip.tmp_h, ip.tmp_w), m2dx(ptr[4][scale], ip.tmp_h, ip.tmp_w), m2dy(ptr[5][scale], ip.tmp_h, ip.tmp_w);
        for (int i = 0; i < n; i += (n / 12 > 0 ? n / 12 : 1)) {
          Patch p(&ip, &fp);
          p.init_patch(&m1, &mdx, &mdy, grid.get_patch_ref_pos(i));
          p.set_target_image(&m2, &m2dx, &m2dy);
          p.inverse_search(Vector2f(0.f, 0.f));
          Vector2f* u = p.get_u_vector();
          const Vector2f q = grid.get_patch_query_pos(i), r = grid.get_patch_ref_pos(i);
          if (!p.is_valid() || p.get_patch_pos()[0] != q[0] || p.get_patch_pos()[1] != q[1] || q[0] != r[0] + (*u)[0]) {
            std::fprintf(stderr, "Patch and PatchGrid disagree on patch %d\n", i);
            return 4;
          }
          patch_id.push_back(i); patch_u.push_back((*u)[0]); patch_u.push_back((*u)[1]);
        }
      }
    }
    FILE* o = std::fopen(argv[2], "wb");
    if (!o) return 2;
    std::fwrite(flow_ctor.data(), 4, flow_ctor.size(), o);
    std::fwrite(flows[finest].data(), 4, flows[finest].size(), o);
    const int K = (int)patch_id.size();
    std::fwrite(&K, 4, 1, o);
    for (int i = 0; i < K; ++i) { std::fwrite(&patch_id[i], 4, 1, o); std::fwrite(&patch_u[2 * i], 4, 2, o); }
    std::fclose(o);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "%s\n", e.what());
    return 3;
  }
  return 0;
}
