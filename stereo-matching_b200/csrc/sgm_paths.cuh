// sgm_paths.cuh -- closed form of the reference SGM path set, and the lane<->disparity map.
//
// The reference enumerates H + W + 2(W+H-1) descriptors on the host every call
// (SGPixelPath::GeneratePaths, reference src/semiglobal.cpp:18-91) and runs each one forward
// and as inverse() (src/semiglobal.hpp:42-45).  For the diagonal descriptors whose true end
// lies on the bottom edge the stored `end` is the image corner, so inverse() restarts
// from the corner: some up-diagonals are visited 0 times and the corner diagonal many
// times.  The multiset of DIRECTED paths is therefore (verified against a
// literal enumeration by tests/test_oracle_golden.py::test_path_set_multiplicities and by
// the GPU parity tests):
//   dir 0 (+1, 0)  H paths, dir 1 (-1, 0)  H paths            multiplicity 1
//   dir 2 ( 0,+1)  W paths, dir 3 ( 0,-1)  W paths            multiplicity 1
//   dir 4 (+1,+1)  W+H-1 diagonals  c = x - y                 multiplicity 1
//   dir 5 (-1,-1)  same diagonals, c* = W - H:
//        W > H : mult 0 for 0 <= c < c*,  c*+1 for c = c*, else 1
//        H > W : mult 0 for c* < c <= -1, -c*  for c = c*, else 1
//   dir 6 (-1,+1)  W+H-1 anti-diagonals s = x + y             multiplicity 1
//   dir 7 (+1,-1)  same anti-diagonals:
//        W > H : mult 0 for H <= s <= W-1, W-H+1 for s = H-1, else 1
// No table is built or uploaded: every kernel derives its path from (dir, index).
#pragma once

#include "stm_common.cuh"

namespace stm {

struct PathGeom {
  int sx, sy;  // first pixel
  int dx, dy;  // step
  int n;       // pixels on the path
  int mult;    // how many times the reference runs this directed path (0 = never)
};

__host__ __device__ inline int sgm_num_paths(int dir, int H, int W) {
  return dir < 2 ? H : (dir < 4 ? W : W + H - 1);
}

__host__ __device__ inline PathGeom sgm_path_geom(int dir, int idx, int H, int W) {
  PathGeom g;
  g.mult = 1;
  switch (dir) {
    case 0: g.sx = 0;     g.sy = idx;   g.dx = 1;  g.dy = 0;  g.n = W; break;
    case 1: g.sx = W - 1; g.sy = idx;   g.dx = -1; g.dy = 0;  g.n = W; break;
    case 2: g.sx = idx;   g.sy = 0;     g.dx = 0;  g.dy = 1;  g.n = H; break;
    case 3: g.sx = idx;   g.sy = H - 1; g.dx = 0;  g.dy = -1; g.n = H; break;
    case 4:
    case 5: {
      const int c = idx - (H - 1);
      const int x0 = c > 0 ? c : 0, y0 = c < 0 ? -c : 0;
      const int n = (W - x0) < (H - y0) ? (W - x0) : (H - y0);
      g.n = n;
      if (dir == 4) {
        g.sx = x0; g.sy = y0; g.dx = 1; g.dy = 1;
      } else {
        g.sx = x0 + n - 1; g.sy = y0 + n - 1; g.dx = -1; g.dy = -1;
        const int cs = W - H;
        if (W > H) {
          if (c >= 0 && c < cs) g.mult = 0;
          else if (c == cs) g.mult = cs + 1;
        } else if (H > W) {
          if (c > cs && c <= -1) g.mult = 0;
          else if (c == cs) g.mult = -cs;
        }
      }
      break;
    }
    default: {
      const int s = idx;
      const int x0 = s < W - 1 ? s : W - 1, y0 = s > W - 1 ? s - (W - 1) : 0;
      const int n = (x0 + 1) < (H - y0) ? (x0 + 1) : (H - y0);
      g.n = n;
      if (dir == 6) {
        g.sx = x0; g.sy = y0; g.dx = -1; g.dy = 1;
      } else {
        g.sx = x0 - (n - 1); g.sy = y0 + n - 1; g.dx = 1; g.dy = -1;
        if (W > H) {
          if (s >= H && s <= W - 1) g.mult = 0;
          else if (s == H - 1) g.mult = W - H + 1;
        }
      }
      break;
    }
  }
  return g;
}

// Multiplicity of the dir-5 / dir-7 path through pixel (x, y): the table above per pixel.
__host__ __device__ inline int sgm_mult5(int x, int y, int H, int W) {
  const int c = x - y, cs = W - H;
  if (W > H) return (c >= 0 && c < cs) ? 0 : (c == cs ? cs + 1 : 1);
  if (H > W) return (c > cs && c <= -1) ? 0 : (c == cs ? -cs : 1);
  return 1;
}
__host__ __device__ inline int sgm_mult7(int x, int y, int H, int W) {
  const int s = x + y;
  if (W > H) return (s >= H && s <= W - 1) ? 0 : (s == H - 1 ? W - H + 1 : 1);
  return 1;
}

// Index of the one path with multiplicity > 1 in dir 5 / dir 7, or -1.
__host__ __device__ inline int sgm_multi_path_index(int dir, int H, int W) {
  if (dir == 5 && W != H) return (W - H) + (H - 1);
  if (dir == 7 && W > H) return H - 1;
  return -1;
}

// ---- lane <-> disparity map ----------------------------------------------------------
// A warp owns one pixel's D-vector.  Lane l holds NV vectors of VW consecutive
// disparities: element (k, j) is d = k*32*VW + l*VW + j, so each of the NV loads is one
// fully coalesced 128*VW-byte warp access.  VW = 4 when D % 128 == 0, 2 when D % 64 == 0,
// else 1 with the tail (d >= D) predicated off and held at +inf (which is exactly the
// reference's out-of-range neighbour value, src/bordered_buffer.hpp:13-18).
struct LaneMap {
  int NV, VW;
};
inline LaneMap pick_lane_map(int D) {
  LaneMap m;
  if (D % 128 == 0) { m.VW = 4; m.NV = D / 128; }
  else if (D % 64 == 0) { m.VW = 2; m.NV = D / 64; }
  else { m.VW = 1; m.NV = (D + 31) / 32; }
  return m;
}

}  // namespace stm
