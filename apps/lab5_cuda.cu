// lab5 contract binaries (designed here: the reference has data files only).
//   cuda_exe_{int,float,uchar}  binary stdin  "int32 n, n keys"  -> binary stdout: n sorted keys
//   to_plot_cuda_exe            text stdin "k1 k2 in_path out_path [int|float|uchar|uint]"
//                               prints the time line; out file = "int32 n, n sorted keys"
// k1/k2 are consumed for harness compatibility; the radix sort picks its own geometry.
// The sort is timed cold, one call (histogram + scan + passes), like the other labs'
// kernels; the file-based build moves large arrays through the pipelined path of
// cli_pipeline.h.
#include "cli_pipeline.h"
#include "lab5_common.h"

#ifndef KEY_TYPE
#define KEY_TYPE 0
#endif

struct DeviceSort {
  void *d_in = nullptr, *d_keys = nullptr, *d_tmp = nullptr, *d_scratch = nullptr;
  size_t bytes = 0, sbytes = 0;
  int32_t n = 0;
  int kt = 0;
  void alloc(int32_t n_, int kt_) {
    n = n_, kt = kt_;
    bytes = (size_t)n * key_size(kt);
    sbytes = b200_sort_scratch_bytes((size_t)n, kt);
    CSC(cudaMalloc(&d_keys, bytes ? bytes : 4));
    CSC(cudaMalloc(&d_tmp, bytes ? bytes : 4));
    CSC(cudaMalloc(&d_scratch, sbytes ? sbytes : 4));
    LIBOK(b200_preload_lab5(kt));
  }
  // the keys are in d_keys; sorted in place
  float run() {
    if (n == 0) return 0.f;
    return cli::time_device_ms(
        [&] { LIBOK(b200_sort(d_keys, d_tmp, (size_t)n, kt, d_scratch, sbytes, nullptr)); },
        /*allow_warmup=*/false);  // in place: a second run would see sorted keys
  }
  void release() { cudaFree(d_keys), cudaFree(d_tmp), cudaFree(d_scratch); }
};

int main() {
#ifdef TO_PLOT
  cli::Input in;
  in.slurp(stdin);
  int k1, k2;
  std::string in_path, out_path, type_word;
  if (!in.next_int(k1) || !in.next_int(k2)) cli::die("expected launch dims: k1 k2");
  if (!in.next_word(in_path) || !in.next_word(out_path)) cli::die("expected input and output paths");
  int kt = key_type_from_name(in_path.substr(in_path.find_last_of('/') + 1), B200_KEY_I32);
  if (in.next_word(type_word)) kt = key_type_from_name(type_word, kt);
  const int in_fd = open(in_path.c_str(), O_RDONLY);
  if (in_fd < 0) cli::die("cannot open input array '%s'", in_path.c_str());
  int32_t n = 0;
  if (pread(in_fd, &n, 4, 0) != 4) cli::die("%s: missing the int32 element count", in_path.c_str());
  if (n < 0) cli::die("%s: negative element count", in_path.c_str());
  struct stat sb;
  const size_t need = (size_t)n * key_size(kt);
  if (fstat(in_fd, &sb) != 0 || (size_t)sb.st_size < 4 + need)
    cli::die("%s: fewer elements than the header promises", in_path.c_str());
  const bool piped = need >= cli::kPipeMin && !cli::same_file(in_path, out_path);
  cli::OutputFile out;
  if (piped) out.begin(out_path, &n, 4, 4 + need);
  DeviceSort ds;
  ds.alloc(n, kt);
  float ms = 0;
  if (piped) {
    cli::PinnedRing ring;
    cli::file_to_device(in_fd, 4, ds.d_keys, need, ring, in_path.c_str());
    ms = ds.run();
    printf("CUDA execution time: <%f ms>\n", ms);
    out.ready();
    cli::device_to_file(out.fd, 4, ds.d_keys, need, ring, out_path.c_str());
    out.finish();
  } else {
    std::vector<uint8_t> keys(need);
    if (need) cli::pread_full(in_fd, keys.data(), need, 4, in_path.c_str());
    CSC(cudaMemcpy(ds.d_keys, keys.data(), need, cudaMemcpyHostToDevice));
    ms = ds.run();
    CSC(cudaMemcpy(keys.data(), ds.d_keys, need, cudaMemcpyDeviceToHost));
    printf("CUDA execution time: <%f ms>\n", ms);
    FILE* o = fopen(out_path.c_str(), "wb");
    if (!o) cli::die("cannot open output array '%s'", out_path.c_str());
    if (fwrite(&n, 4, 1, o) != 1 || fwrite(keys.data(), 1, keys.size(), o) != keys.size())
      cli::die("short write to '%s'", out_path.c_str());
    fclose(o);
  }
  close(in_fd);
  ds.release();
#else
  const int kt = KEY_TYPE;
  std::vector<uint8_t> raw = read_all(stdin);
  int32_t n;
  std::vector<uint8_t> keys = parse_array(raw, kt, n, "stdin");
  DeviceSort ds;
  ds.alloc(n, kt);
  CSC(cudaMemcpy(ds.d_keys, keys.data(), keys.size(), cudaMemcpyHostToDevice));
  ds.run();
  CSC(cudaMemcpy(keys.data(), ds.d_keys, keys.size(), cudaMemcpyDeviceToHost));
  ds.release();
  fwrite(keys.data(), 1, keys.size(), stdout);
#endif
  return 0;
}
