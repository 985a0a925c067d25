// lab3 CPU binary (new: the reference has no CPU build of lab3).  Same stdin as
// lab3/src/cuda_exe, prints "CPU execution time: <t ms>".  Statistics follow
// B200LABS_LAB3_STATS like the CUDA build (default: integer moments).
#include <chrono>

#include "lab3_common.h"

int main() {
  cli::Input in;
  in.slurp(stdin);
  Lab3Spec spec = read_lab3_spec(in);
  cli::Image im = cli::read_image(spec.in_path);
  check_points(spec, im.w, im.h);
  std::vector<double> params((size_t)spec.nc * B200_LAB3_PARAMS);
  const char* mode = getenv("B200LABS_LAB3_STATS");
  if (mode && !strcmp(mode, "host")) {
    b200_cpu_lab3_stats_ref(im.px.data(), im.w, spec.pts.data(), spec.class_off.data(), spec.nc,
                            params.data());
  } else {
    std::vector<long long> mom((size_t)spec.nc * B200_LAB3_MOMENTS, 0);
    b200_cpu_lab3_moments(im.px.data(), im.w, 0, im.h, spec.pts.data(), spec.class_off.data(),
                          spec.nc, mom.data());
    b200_cpu_lab3_finalize(mom.data(), spec.nc, params.data());
  }
  auto t0 = std::chrono::steady_clock::now();
  b200_cpu_lab3_classify(im.px.data(), (int64_t)im.w * im.h, spec.nc, params.data());
  auto t1 = std::chrono::steady_clock::now();
  printf("CPU execution time: <%f ms>\n", std::chrono::duration<double, std::milli>(t1 - t0).count());
  cli::write_image(spec.out_path, im.w, im.h, im.px.data());
  return 0;
}
