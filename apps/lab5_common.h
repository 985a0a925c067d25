// lab5 array files: little-endian int32 n, then n keys of the build's type
// (layout of lab5/data/{int10,float10,uchar10} in the reference).
#pragma once
#include "b200labs.h"
#include "cli_common.h"

inline size_t key_size(int kt) { return kt == B200_KEY_U8 ? 1 : 4; }

inline int key_type_from_name(const std::string& s, int fallback) {
  if (s.find("float") != std::string::npos) return B200_KEY_F32;
  if (s.find("uchar") != std::string::npos) return B200_KEY_U8;
  if (s.find("uint") != std::string::npos) return B200_KEY_U32;
  if (s.find("int") != std::string::npos) return B200_KEY_I32;
  return fallback;
}

inline std::vector<uint8_t> read_all(FILE* f) {
  std::vector<uint8_t> buf;
  uint8_t chunk[1 << 16];
  size_t got;
  while ((got = fread(chunk, 1, sizeof chunk, f)) > 0) buf.insert(buf.end(), chunk, chunk + got);
  return buf;
}

// returns the key bytes; n comes back through the reference
inline std::vector<uint8_t> parse_array(const std::vector<uint8_t>& raw, int kt, int32_t& n,
                                        const char* what) {
  if (raw.size() < 4) cli::die("%s: missing the int32 element count", what);
  memcpy(&n, raw.data(), 4);
  if (n < 0) cli::die("%s: negative element count", what);
  size_t need = (size_t)n * key_size(kt);
  if (raw.size() - 4 < need) cli::die("%s: fewer elements than the header promises", what);
  return std::vector<uint8_t>(raw.begin() + 4, raw.begin() + 4 + need);
}
