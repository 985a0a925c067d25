// lab2 CPU binary: stdin "in out", prints "CPU execution time: <t ms>"
// (ref: lab2/src/main.c:64-65,106).  Arithmetic is the FMA-faithful oracle, so
// unlike the reference's gcc build it is bit-identical to the GPU path on every
// image, not only on the golden files.
#include <chrono>

#include "b200labs.h"
#include "cli_common.h"

int main() {
  cli::Input in;
  in.slurp(stdin);
  std::string in_path, out_path;
  if (!in.next_word(in_path) || !in.next_word(out_path)) cli::die("expected input and output paths");
  cli::Image im = cli::read_image(in_path);
  std::vector<uint32_t> out(im.px.size());
  auto t0 = std::chrono::steady_clock::now();
  b200_cpu_roberts(im.px.data(), out.data(), im.w, im.h);
  auto t1 = std::chrono::steady_clock::now();
  printf("CPU execution time: <%f ms>\n", std::chrono::duration<double, std::milli>(t1 - t0).count());
  cli::write_image(out_path, im.w, im.h, out.data());
  return 0;
}
