// lab3 stdin protocol: "in out nc" then per class "np x1 y1 ... x_np y_np"
// (ref: lab3/src/main.cu:81-82,97-109; limits nc <= 32, np <= 2^19).
#pragma once
#include "b200labs.h"
#include "cli_common.h"

struct Lab3Spec {
  std::string in_path, out_path;
  int nc = 0;
  std::vector<int> pts;        // x,y pairs, all classes back to back
  std::vector<int> class_off;  // nc + 1 offsets into pts (in points)
};

inline Lab3Spec read_lab3_spec(cli::Input& in) {
  Lab3Spec s;
  if (!in.next_word(s.in_path) || !in.next_word(s.out_path)) cli::die("expected input and output paths");
  if (!in.next_int(s.nc) || s.nc < 1 || s.nc > B200_LAB3_MAX_CLASSES)
    cli::die("class count must be 1..32");
  s.class_off.push_back(0);
  for (int c = 0; c < s.nc; ++c) {
    int np = 0;
    if (!in.next_int(np) || np < 0) cli::die("expected the number of sample points of a class");
    for (int i = 0; i < 2 * np; ++i) {
      int v;
      if (!in.next_int(v)) cli::die("class sample list is shorter than its count");
      s.pts.push_back(v);
    }
    s.class_off.push_back(s.class_off.back() + np);
  }
  return s;
}

inline void check_points(const Lab3Spec& s, int w, int h) {
  for (size_t i = 0; i + 1 < s.pts.size(); i += 2)
    if (s.pts[i] < 0 || s.pts[i] >= w || s.pts[i + 1] < 0 || s.pts[i + 1] >= h)
      cli::die("a sample point lies outside the image");
}
