// lab5 CPU binaries: same protocols as the CUDA builds, std::sort as the timed
// CPU baseline and oracle.  to_plot_cpu_exe takes no launch-dims prefix
// ("in_path out_path [type]"), like the other labs' CPU builds.
#include <chrono>

#include "lab5_common.h"

#ifndef KEY_TYPE
#define KEY_TYPE 0
#endif

int main() {
#ifdef TO_PLOT
  cli::Input in;
  in.slurp(stdin);
  std::string in_path, out_path, type_word;
  if (!in.next_word(in_path) || !in.next_word(out_path)) cli::die("expected input and output paths");
  int kt = key_type_from_name(in_path.substr(in_path.find_last_of('/') + 1), B200_KEY_I32);
  if (in.next_word(type_word)) kt = key_type_from_name(type_word, kt);
  FILE* f = fopen(in_path.c_str(), "rb");
  if (!f) cli::die("cannot open input array '%s'", in_path.c_str());
  std::vector<uint8_t> raw = read_all(f);
  fclose(f);
  int32_t n;
  std::vector<uint8_t> keys = parse_array(raw, kt, n, in_path.c_str());
  auto t0 = std::chrono::steady_clock::now();
  b200_cpu_sort(keys.data(), (size_t)n, kt);
  auto t1 = std::chrono::steady_clock::now();
  printf("CPU execution time: <%f ms>\n", std::chrono::duration<double, std::milli>(t1 - t0).count());
  FILE* o = fopen(out_path.c_str(), "wb");
  if (!o) cli::die("cannot open output array '%s'", out_path.c_str());
  fwrite(&n, 4, 1, o);
  fwrite(keys.data(), 1, keys.size(), o);
  fclose(o);
#else
  const int kt = KEY_TYPE;
  std::vector<uint8_t> raw = read_all(stdin);
  int32_t n;
  std::vector<uint8_t> keys = parse_array(raw, kt, n, "stdin");
  b200_cpu_sort(keys.data(), (size_t)n, kt);
  fwrite(keys.data(), 1, keys.size(), stdout);
#endif
  return 0;
}
