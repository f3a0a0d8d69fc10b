// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load this library.
//
// CPU restatement ("oracle") of the reference's ORB front-end hot path:
//   ORBextractor  (ctor tables, pyramid, per-cell FAST, DistributeOctTree, IC_Angle,
//                  blur, steered BRIEF, output packing)   /root/reference/Features/ORBextractor.cpp
//   Frame glue    (undistort, image bounds, 64x48 grid, GetFeaturesInArea)
//                                                      /root/reference/SlamTypes/Frame.cpp
//   ORBmatcher::SearchForInitialization + FORB::distance
//                                                      /root/reference/Features/ORBmatcher.cpp,
//                                                      /root/reference/Thirdparty/DBoW2/src/FORB.cpp
// Each function cites the reference lines it follows.  The OpenCV primitives the reference
// calls are restated in cv_prims.hpp and pinned against cv2 4.13.0 by the tests.
//
// Parity status: the reference has no tests, golden vectors or fixtures, so
// this oracle is pinned (a) per OpenCV primitive against the real OpenCV 4.13.0 wheel and
// (b) for the reference's own logic against oracle/_ref (the reference's own .cpp files
// compiled over a header shim, see oracle/build_ref.sh) whenever that library is present.
//
// Semantics that matter and are kept on purpose:
//  * std::list + std::sort (libstdc++ introsort tie order) in the quadtree, as in the reference;
//  * float/double mixes of the constructor exactly as written (member scaleFactor is double);
//  * cos/sin of a float angle evaluated in double and narrowed (unqualified ::cos(double),
//    Features/ORBextractor.cpp:174); a cosf/sinf variant is available for the mismatch report;
//  * x86-64 baseline arithmetic: no FMA, round-half-even cvRound.
#include <algorithm>
#include <atomic>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <list>
#include <thread>
#include <utility>
#include <vector>

#include "cv_prims.hpp"

namespace orc {

static const int kPatch = 31, kHalfPatch = 15, kEdge = 19;

static const int kPatternTable[256 * 4] = {
#include "../orb_slam_tracking_b200/csrc/orb_pattern.inc"
};

struct KeyPoint {  // layout-compatible with cv::KeyPoint (28 bytes)
  float x, y, size, angle, response;
  int octave, class_id;
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

// =====================================================================================
// Extractor tables -- ORBextractor::ORBextractor, Features/ORBextractor.cpp:492-595
// =====================================================================================
struct Tables {
  int nfeatures, nlevels, iniTh, minTh;
  double scaleFactorD;  // the member is `double scaleFactor` holding a float (hpp:142)
  std::vector<float> scale, invScale, sigma2, invSigma2;
  std::vector<int> perLevel, umax;
  int status;  // 0 ok, -1: reference would exit(1)
};

static Tables make_tables(int nfeatures, float scaleFactorF, int nlevels, int iniTh, int minTh) {
  Tables t;
  t.nfeatures = nfeatures; t.nlevels = nlevels; t.iniTh = iniTh; t.minTh = minTh;
  t.scaleFactorD = scaleFactorF;
  t.status = 0;
  if (t.scaleFactorD == 1.0 && nlevels > 1) { t.status = -1; return t; }  // :502-505
  t.scale.assign(nlevels, 1.0f);
  t.sigma2.assign(nlevels, 1.0f);
  for (int i = 1; i < nlevels; ++i) {
    t.scale[i] = (float)(t.scale[i - 1] * t.scaleFactorD);  // float * double -> float, :512
    t.sigma2[i] = t.scale[i] * t.scale[i];                  // :513
  }
  t.invScale.resize(nlevels);
  t.invSigma2.resize(nlevels);
  for (int i = 0; i < nlevels; ++i) {
    t.invScale[i] = 1.0f / t.scale[i];
    t.invSigma2[i] = 1.0f / t.sigma2[i];
  }
  // per-level quotas, :529-548
  t.perLevel.assign(nlevels, 0);
  float factor = (float)(1.0f / t.scaleFactorD);
  float desired = nfeatures * (1 - factor) /
                  (1 - (float)std::pow((double)factor, (double)(float)nlevels));
  int sum = 0;
  for (int l = 0; l < nlevels - 1; ++l) {
    t.perLevel[l] = cv_round(desired);
    sum += t.perLevel[l];
    desired *= factor;
  }
  t.perLevel[nlevels - 1] = std::max(nfeatures - sum, 0);
  // umax circular-patch row extents, :562-594
  t.umax.assign(kHalfPatch + 1, 0);
  int vmax = cv_floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
  int vmin = cv_ceil(kHalfPatch * std::sqrt(2.f) / 2);
  const double hp2 = kHalfPatch * kHalfPatch;
  for (int v = 0; v <= vmax; ++v) t.umax[v] = cv_round(std::sqrt(hp2 - v * v));
  for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
    while (t.umax[v0] == t.umax[v0 + 1]) ++v0;
    t.umax[v] = v0;
    ++v0;
  }
  return t;
}

// =====================================================================================
// Pyramid -- ORBextractor::ComputePyramid, Features/ORBextractor.cpp:1660-1713
// Each level is stored padded by 19 px (BORDER_REFLECT_101); `roi()` is the image proper.
// =====================================================================================
struct Level {
  int w, h;             // unpadded size
  std::vector<u8> pad;  // (w+38) x (h+38)
  size_t step() const { return (size_t)w + 2 * kEdge; }
  const u8* roi() const { return pad.data() + (size_t)kEdge * step() + kEdge; }
  u8* roi() { return pad.data() + (size_t)kEdge * step() + kEdge; }
};

static void build_pyramid(const Tables& t, const u8* img, int W, int H, size_t step,
                          std::vector<Level>& pyr) {
  pyr.resize(t.nlevels);
  for (int l = 0; l < t.nlevels; ++l) {
    float s = t.invScale[l];
    Level& L = pyr[l];
    L.w = cv_round(W * s);  // int * float -> float, :1663
    L.h = cv_round(H * s);
    L.pad.assign(L.step() * (size_t)(L.h + 2 * kEdge), 0);
    if (l == 0) {
      copy_make_border101(img, W, H, step, L.pad.data(), L.step(), kEdge);
    } else {
      const Level& P = pyr[l - 1];
      // source is the unpadded ROI of the previous level (:1676-1682)
      resize_linear_u8(P.roi(), P.w, P.h, P.step(), L.roi(), L.w, L.h, L.step());
      // BORDER_ISOLATED: the border is reflected from the ROI itself (:1689-1695)
      std::vector<u8> tmp((size_t)L.w * L.h);
      for (int y = 0; y < L.h; ++y) std::memcpy(&tmp[(size_t)y * L.w], L.roi() + y * L.step(), L.w);
      copy_make_border101(tmp.data(), L.w, L.h, L.w, L.pad.data(), L.step(), kEdge);
    }
  }
}

// =====================================================================================
// Quadtree distribution -- ExtractorNode::DivideNode :617-676, compareNodes :684-696,
// ORBextractor::DistributeOctTree :698-1011
// =====================================================================================
struct Cand { float x, y, response; };  // the only KeyPoint fields the tree reads

struct QNode {
  std::vector<Cand> keys;
  int ulx, uly, urx, ury, blx, bly, brx, bry;
  std::list<QNode>::iterator self;
  bool leaf;  // bNoMore
  QNode() : ulx(0), uly(0), urx(0), ury(0), blx(0), bly(0), brx(0), bry(0), leaf(false) {}
};

static void divide(const QNode& p, QNode& a, QNode& b, QNode& c, QNode& d) {
  const int hx = (int)std::ceil((float)(p.urx - p.ulx) / 2);
  const int hy = (int)std::ceil((float)(p.bry - p.uly) / 2);
  a.ulx = p.ulx; a.uly = p.uly; a.urx = p.ulx + hx; a.ury = p.uly;
  a.blx = p.ulx; a.bly = p.uly + hy; a.brx = p.ulx + hx; a.bry = p.uly + hy;
  b.ulx = a.urx; b.uly = a.ury; b.urx = p.urx; b.ury = p.ury;
  b.blx = a.brx; b.bly = a.bry; b.brx = p.urx; b.bry = p.uly + hy;
  c.ulx = a.blx; c.uly = a.bly; c.urx = a.brx; c.ury = a.bry;
  c.blx = p.blx; c.bly = p.bly; c.brx = a.brx; c.bry = p.bly;
  d.ulx = c.urx; d.uly = c.ury; d.urx = b.brx; d.ury = b.bry;
  d.blx = c.brx; d.bly = c.bry; d.brx = p.brx; d.bry = p.bry;
  for (size_t i = 0; i < p.keys.size(); ++i) {
    const Cand& k = p.keys[i];
    if (k.x < a.urx) {
      if (k.y < a.bry) a.keys.push_back(k); else c.keys.push_back(k);
    } else if (k.y < a.bry) {
      b.keys.push_back(k);
    } else {
      d.keys.push_back(k);
    }
  }
  a.leaf = a.keys.size() == 1; b.leaf = b.keys.size() == 1;
  c.leaf = c.keys.size() == 1; d.leaf = d.keys.size() == 1;
}

typedef std::pair<int, QNode*> SizedNode;
static bool sized_less(SizedNode& l, SizedNode& r) {  // non-const refs as in the reference
  if (l.first < r.first) return true;
  if (l.first > r.first) return false;
  return l.second->ulx < r.second->ulx;
}

// returns -2 if the reference would index an empty vector (aspect ratio rounds to 0)
static int distribute_quadtree(const std::vector<Cand>& in, int minX, int maxX, int minY, int maxY,
                               int N, std::vector<Cand>& out) {
  out.clear();
  const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));
  if (nIni < 1) return -2;
  const float hX = (float)(maxX - minX) / nIni;

  std::list<QNode> nodes;
  std::vector<QNode*> roots(nIni);
  for (int i = 0; i < nIni; ++i) {
    QNode n;
    n.ulx = (int)(hX * (float)i); n.uly = 0;
    n.urx = (int)(hX * (float)(i + 1)); n.ury = 0;
    n.blx = n.ulx; n.bly = maxY - minY;
    n.brx = n.urx; n.bry = maxY - minY;
    nodes.push_back(n);
    roots[i] = &nodes.back();
  }
  for (size_t i = 0; i < in.size(); ++i) roots[(size_t)(in[i].x / hX)]->keys.push_back(in[i]);

  for (std::list<QNode>::iterator it = nodes.begin(); it != nodes.end();) {
    if (it->keys.size() == 1) { it->leaf = true; ++it; }
    else if (it->keys.empty()) it = nodes.erase(it);
    else ++it;
  }

  std::vector<SizedNode> splittable;
  splittable.reserve(nodes.size() * 4);
  // pushes the non-empty children in the order a,b,c,d; records those with >1 key
  struct Push {
    static int child(std::list<QNode>& L, std::vector<SizedNode>& S, QNode& c) {
      if (c.keys.empty()) return 0;
      L.push_front(c);
      if (c.keys.size() > 1) {
        S.push_back(std::make_pair((int)c.keys.size(), &L.front()));
        L.front().self = L.begin();
        return 1;
      }
      return 0;
    }
  };

  bool done = false;
  while (!done) {
    int prevSize = (int)nodes.size();
    int nToExpand = 0;
    splittable.clear();
    for (std::list<QNode>::iterator it = nodes.begin(); it != nodes.end();) {
      if (it->leaf) { ++it; continue; }
      QNode a, b, c, d;
      divide(*it, a, b, c, d);
      nToExpand += Push::child(nodes, splittable, a);
      nToExpand += Push::child(nodes, splittable, b);
      nToExpand += Push::child(nodes, splittable, c);
      nToExpand += Push::child(nodes, splittable, d);
      it = nodes.erase(it);
    }
    if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
      done = true;
    } else if ((int)nodes.size() + nToExpand * 3 > N) {
      while (!done) {
        prevSize = (int)nodes.size();
        std::vector<SizedNode> prev = splittable;
        splittable.clear();
        std::sort(prev.begin(), prev.end(), sized_less);
        for (int j = (int)prev.size() - 1; j >= 0; --j) {
          QNode a, b, c, d;
          divide(*prev[j].second, a, b, c, d);
          Push::child(nodes, splittable, a);
          Push::child(nodes, splittable, b);
          Push::child(nodes, splittable, c);
          Push::child(nodes, splittable, d);
          nodes.erase(prev[j].second->self);
          if ((int)nodes.size() >= N) break;
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) done = true;
      }
    }
  }
  // first strict maximum of the response per node, list order (:984-1007)
  out.reserve(nodes.size());
  for (std::list<QNode>::iterator it = nodes.begin(); it != nodes.end(); ++it) {
    const Cand* best = &it->keys[0];
    for (size_t k = 1; k < it->keys.size(); ++k)
      if (it->keys[k].response > best->response) best = &it->keys[k];
    out.push_back(*best);
  }
  return 0;
}

// =====================================================================================
// Per-cell FAST -- ComputeKeyPointsOctTree cell loop, Features/ORBextractor.cpp:1051-1141
// Output coordinates are relative to (minBorderX, minBorderY) = (16, 16).
// =====================================================================================
static void detect_cells(const Level& L, int iniTh, int minTh, std::vector<Cand>& out) {
  out.clear();
  const int minBX = kEdge - 3, minBY = minBX;
  const int maxBX = L.w - kEdge + 3, maxBY = L.h - kEdge + 3;
  const float Wc = 35;
  const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
  const int nCols = (int)(width / Wc), nRows = (int)(height / Wc);
  if (nCols <= 0 || nRows <= 0) return;  // reference loops do not execute
  const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
  std::vector<FastPt> cell;
  for (int i = 0; i < nRows; ++i) {
    const float iniY = (float)(minBY + i * hCell);
    float maxY = iniY + hCell + 6;
    if (iniY >= maxBY - 3) continue;
    if (maxY > maxBY) maxY = (float)maxBY;
    for (int j = 0; j < nCols; ++j) {
      const float iniX = (float)(minBX + j * wCell);
      float maxX = iniX + wCell + 6;
      if (iniX >= maxBX - 6) continue;
      if (maxX > maxBX) maxX = (float)maxBX;
      const int x0 = (int)iniX, x1 = (int)maxX, y0 = (int)iniY, y1 = (int)maxY;
      const u8* roi = L.roi() + (size_t)y0 * L.step() + x0;
      fast9_16_nms(roi, x1 - x0, y1 - y0, L.step(), iniTh, cell);
      if (cell.empty()) fast9_16_nms(roi, x1 - x0, y1 - y0, L.step(), minTh, cell);
      for (size_t k = 0; k < cell.size(); ++k) {
        Cand c = {(float)cell[k].x + j * wCell, (float)cell[k].y + i * hCell, (float)cell[k].score};
        out.push_back(c);
      }
    }
  }
}

// =====================================================================================
// ComputeKeyPointsOld, Features/ORBextractor.cpp:1207-1496: the alternate selection policy the
// reference keeps behind a commented-out call (:1555).  Larger cells (about five keypoints per cell wanted), the
// minThFAST retry when a cell yields at most 3 keypoints, a redistribution of the per-cell quotas towards the cells that
// have more, and cv::KeyPointsFilter::retainBest per cell and per level.  Returns -7 where the reference divides by zero
// (a level whose quota is too small for one cell column).
// =====================================================================================
static int keypoints_old(const Level& L, const Tables& t, int level, float imageRatio, std::vector<KeyPoint>& keypoints) {
  keypoints.clear();
  const int nDesired = t.perLevel[level];
  const int levelCols = (int)std::sqrt((float)nDesired / (5 * imageRatio));  // :1225-1226
  const int levelRows = (int)(imageRatio * levelCols);                       // :1228
  if (levelCols < 1 || levelRows < 1) return -7;
  const int minBX = kEdge, minBY = kEdge, maxBX = L.w - kEdge, maxBY = L.h - kEdge;  // :1233-1236
  const int W = maxBX - minBX, H = maxBY - minBY;
  const int cellW = (int)std::ceil((float)W / levelCols), cellH = (int)std::ceil((float)H / levelRows);
  // Outside the function's working range: a cell row / column other than the last ends at minB + (i + 1) * cell + 3 (:1279, :1309); when
  // that lies beyond maxB + 3 the reference tests and keeps keypoints closer than 19 px to the level's border -- its descriptor then
  // reads outside the blurred clone (:1598-1613, undefined values), and a cell beyond the image makes cv::Mat::rowRange assert.
  // Reported as -8 (the CUDA path refuses the shape with ORBX_ERR_ARG).
  if ((levelRows - 1) * cellH > H || (levelCols - 1) * cellW > W) return -8;
  const int nCells = levelRows * levelCols;
  const int nfeaturesCell = (int)std::ceil((float)nDesired / nCells);  // :1250-1251
  std::vector<std::vector<KeyPoint>> cellKeys((size_t)nCells);
  std::vector<int> nToRetain(nCells, 0), nTotal(nCells, 0), iniXCol(levelCols), iniYRow(levelRows);
  std::vector<char> bNoMore(nCells, 0);
  int nNoMore = 0, nToDistribute = 0;
  std::vector<FastPt> cell;
  float hY = (float)(cellH + 6);
  for (int i = 0; i < levelRows; ++i) {
    const float iniY = (float)(minBY + i * cellH - 3);
    iniYRow[i] = (int)iniY;
    if (i == levelRows - 1) {
      hY = maxBY + 3 - iniY;
      if (hY <= 0) continue;
    }
    float hX = (float)(cellW + 6);
    for (int j = 0; j < levelCols; ++j) {
      float iniX;
      if (i == 0) { iniX = (float)(minBX + j * cellW - 3); iniXCol[j] = (int)iniX; }
      else iniX = (float)iniXCol[j];
      if (j == levelCols - 1) {
        hX = maxBX + 3 - iniX;
        if (hX <= 0) continue;
      }
      const int x0 = (int)iniX, y0 = (int)iniY, x1 = (int)(iniX + hX), y1 = (int)(iniY + hY);
      const u8* roi = L.roi() + (size_t)y0 * L.step() + x0;
      fast9_16_nms(roi, x1 - x0, y1 - y0, L.step(), t.iniTh, cell);                       // :1336-1340
      if (cell.size() <= 3) fast9_16_nms(roi, x1 - x0, y1 - y0, L.step(), t.minTh, cell);  // :1343-1351 (fills `cell` afresh)
      std::vector<KeyPoint>& out = cellKeys[(size_t)i * levelCols + j];
      out.resize(cell.size());
      for (size_t k = 0; k < cell.size(); ++k) {
        KeyPoint kp;
        kp.x = (float)cell[k].x; kp.y = (float)cell[k].y; kp.size = 7.f; kp.angle = -1.f; kp.response = (float)cell[k].score;
        kp.octave = 0; kp.class_id = -1;
        out[k] = kp;
      }
      const int c = i * levelCols + j, nKeys = (int)cell.size();
      nTotal[c] = nKeys;
      if (nKeys > nfeaturesCell) { nToRetain[c] = nfeaturesCell; bNoMore[c] = 0; }
      else { nToRetain[c] = nKeys; nToDistribute += nfeaturesCell - nKeys; bNoMore[c] = 1; ++nNoMore; }
    }
  }
  while (nToDistribute > 0 && nNoMore < nCells) {  // :1390-1425
    const int nNew = nfeaturesCell + (int)std::ceil((float)nToDistribute / (nCells - nNoMore));
    nToDistribute = 0;
    for (int c = 0; c < nCells; ++c) {
      if (bNoMore[c]) continue;
      if (nTotal[c] > nNew) { nToRetain[c] = nNew; bNoMore[c] = 0; }
      else { nToRetain[c] = nTotal[c]; nToDistribute += nNew - nTotal[c]; bNoMore[c] = 1; ++nNoMore; }
    }
  }
  const int scaledPatch = (int)(kPatch * t.scale[level]);
  for (int i = 0; i < levelRows; ++i)
    for (int j = 0; j < levelCols; ++j) {
      std::vector<KeyPoint>& kc = cellKeys[(size_t)i * levelCols + j];
      const int keep = nToRetain[i * levelCols + j];
      retain_best(kc, keep);                      // :1452
      if ((int)kc.size() > keep) kc.resize(keep);  // :1455-1456
      for (size_t k = 0; k < kc.size(); ++k) {
        kc[k].x += iniXCol[j]; kc[k].y += iniYRow[i];
        kc[k].octave = level; kc[k].size = (float)scaledPatch;
        keypoints.push_back(kc[k]);
      }
    }
  if ((int)keypoints.size() > nDesired) {  // :1480-1484
    retain_best(keypoints, nDesired);
    keypoints.resize(nDesired);
  }
  return 0;
}

// =====================================================================================
// Orientation -- IC_Angle, Features/ORBextractor.cpp:103-159
// =====================================================================================
static float ic_angle(const u8* img, size_t step, float px, float py, const std::vector<int>& umax) {
  int m01 = 0, m10 = 0;
  const u8* c = img + (ptrdiff_t)cv_round(py) * (ptrdiff_t)step + cv_round(px);
  for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * c[u];
  const ptrdiff_t s = (ptrdiff_t)step;
  for (int v = 1; v <= kHalfPatch; ++v) {
    int vsum = 0, d = umax[v];
    for (int u = -d; u <= d; ++u) {
      int below = c[v * s + u], above = c[-v * s + u];
      vsum += below - above;
      m10 += u * (below + above);
    }
    m01 += v * vsum;
  }
  return fast_atan2((float)m01, (float)m10);
}

// =====================================================================================
// Steered BRIEF -- computeOrbDescriptor, Features/ORBextractor.cpp:169-228
// `cos(angle)` / `sin(angle)` at :174 take a float under <cmath> + `using namespace std`: the float overloads, i.e. glibc's
// cosf / sinf.  glibc is not part of /root/reference, so its published algorithm is restated here (glibc >= 2.28,
// sysdeps/ieee754/flt-32/{s_sincosf.h,s_sinf.c,s_cosf.c}: the ARM optimized-routines sincosf -- reduction by one multiply with
// 2/pi * 2^24, a degree-7 sine / degree-8 cosine polynomial in double, one rounding to float) and pinned against the libm of
// the machine: tools/probe/glibc_sincosf_check.c compares it on EVERY float of [0, 2 pi] (and up to 119.9), bit for bit;
// tests/test_trig_port.py repeats a sample.  On x86-64 the ifunc-selected __cosf_fma / __sinf_fma contract every a * b + c;
// in the angle range the contracted and the plain form round to the same float everywhere (same check).
// trigMode 2: this restatement [default, pinned]   0: (float)cos((double)a)   1: the machine's cosf(a) / sinf(a)
// =====================================================================================
static inline double fma_d(double a, double b, double c) { return __builtin_fma(a, b, c); }
static void glibc_sincosf_port(float y, float* sn, float* cs) {
  static const double C[5] = {0x1p0, -0x1.ffffffd0c621cp-2, 0x1.55553e1068f19p-5, -0x1.6c087e89a359dp-10, 0x1.99343027bf8c3p-16};
  static const double S[3] = {-0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13};
  uint32_t bits;
  std::memcpy(&bits, &y, 4);
  const uint32_t top = (bits >> 20) & 0x7ff;               // abstop12
  if (top >= 0x42f) { *sn = sinf(y); *cs = cosf(y); return; }  // |y| >= 120, inf, nan: not restated (an angle in radians is <= 2 pi)
  if (top < 0x398) { *sn = y; *cs = 1.0f; return; }   // |y| < 2^-12
  double x = y;
  int n = 0;
  if (top >= 0x3f4) {                                 // |y| >= pi / 4 (float): reduce_fast
    const double r = x * 0x1.45F306DC9C883p+23;
    n = ((int32_t)r + 0x800000) >> 24;
    x = fma_d(-(double)n, 0x1.921FB54442D18p0, x);
  }
  const double sg = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0, ng = (n & 2) ? -1.0 : 1.0;
  const double xs = x * sg, x2 = x * x;
  const double x3 = xs * x2, s1 = fma_d(x2, S[2], S[1]), x7 = x3 * x2, s = fma_d(x3, S[0], xs);
  const float sp = (float)fma_d(x7, s1, s);           // sinf_poly, even quadrant
  const double x4 = x2 * x2, c2 = fma_d(x2, ng * C[4], ng * C[3]), c1 = fma_d(x2, ng * C[1], ng * C[0]), x6 = x4 * x2;
  const double c = fma_d(x4, ng * C[2], c1);
  const float cp = (float)fma_d(x6, c2, c);           // sinf_poly, odd quadrant (second table when n & 2)
  if (n & 1) { *sn = cp; *cs = sp; } else { *sn = sp; *cs = cp; }
}
static void brief256(const u8* img, size_t step, const KeyPoint& kp, int trigMode, u8* desc) {
  const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
  float angle = kp.angle * factorPI;
  float co, si;
  if (trigMode == 0) { co = (float)std::cos((double)angle); si = (float)std::sin((double)angle); }
  else if (trigMode == 1) { co = cosf(angle); si = sinf(angle); }
  else glibc_sincosf_port(angle, &si, &co);
  const u8* c = img + (ptrdiff_t)cv_round(kp.y) * (ptrdiff_t)step + cv_round(kp.x);
  const int* p = kPatternTable;
  for (int i = 0; i < 32; ++i) {
    unsigned v = 0;
    for (int b = 0; b < 8; ++b, p += 4) {
      int r0 = cv_round(p[0] * si + p[1] * co), c0 = cv_round(p[0] * co - p[1] * si);
      int r1 = cv_round(p[2] * si + p[3] * co), c1 = cv_round(p[2] * co - p[3] * si);
      int t0 = c[(ptrdiff_t)r0 * (ptrdiff_t)step + c0], t1 = c[(ptrdiff_t)r1 * (ptrdiff_t)step + c1];
      v |= (unsigned)(t0 < t1) << b;
    }
    desc[i] = (u8)v;
  }
}

// =====================================================================================
// Extractor -- ORBextractor::operator(), :1531-1653; ComputeKeyPointsOctTree :1026-1189
// Keeps every intermediate so tests can compare stage by stage.
// =====================================================================================
struct Extractor {
  Tables t;
  int trigMode;
  int policy;  // 0: ComputeKeyPointsOctTree (what operator() calls, :1552); 1: ComputeKeyPointsOld (:1555, commented out)
  std::vector<Level> pyr;
  std::vector<std::vector<Cand>> cands;        // per level, FAST candidates (16,16)-relative
  std::vector<std::vector<KeyPoint>> levelKps;  // per level, after quadtree + orientation
  std::vector<std::vector<u8>> blurred;          // per level, w*h (empty if level had no kps)
  std::vector<std::vector<u8>> levelDesc;        // per level, n*32
};

static int extract(Extractor& E, const u8* img, int W, int H, size_t step, const int lapping[2],
                   std::vector<KeyPoint>& kps, std::vector<u8>& desc, int* monoIndex) {
  const Tables& t = E.t;
  if (!img || W <= 0 || H <= 0) return -1;  // :1536
  build_pyramid(t, img, W, H, step, E.pyr);
  E.cands.assign(t.nlevels, std::vector<Cand>());
  E.levelKps.assign(t.nlevels, std::vector<KeyPoint>());
  E.blurred.assign(t.nlevels, std::vector<u8>());
  E.levelDesc.assign(t.nlevels, std::vector<u8>());
  for (int l = 0; l < t.nlevels && E.policy == 1; ++l) {
    const int rc = keypoints_old(E.pyr[l], t, l, (float)E.pyr[0].w / E.pyr[0].h, E.levelKps[l]);  // :1213-1214
    if (rc) return rc;
  }
  for (int l = 0; l < t.nlevels && E.policy == 0; ++l) {
    const Level& L = E.pyr[l];
    detect_cells(L, t.iniTh, t.minTh, E.cands[l]);
    const int minBX = kEdge - 3, minBY = minBX;
    const int maxBX = L.w - kEdge + 3, maxBY = L.h - kEdge + 3;
    std::vector<Cand> kept;
    int rc = distribute_quadtree(E.cands[l], minBX, maxBX, minBY, maxBY, t.perLevel[l], kept);
    if (rc) return rc;
    if ((int)kept.size() > t.perLevel[l]) kept.resize(t.perLevel[l]);  // :1159-1161
    const int scaledPatch = (int)(kPatch * t.scale[l]);                 // :1165
    std::vector<KeyPoint>& out = E.levelKps[l];
    out.resize(kept.size());
    for (size_t i = 0; i < kept.size(); ++i) {
      KeyPoint k;
      k.x = kept[i].x + minBX; k.y = kept[i].y + minBY;
      k.size = (float)scaledPatch; k.angle = -1.f; k.response = kept[i].response;
      k.octave = l; k.class_id = -1;
      out[i] = k;
    }
  }
  for (int l = 0; l < t.nlevels; ++l)  // :1184-1188, on the unblurred level
    for (size_t i = 0; i < E.levelKps[l].size(); ++i) {
      KeyPoint& k = E.levelKps[l][i];
      k.angle = ic_angle(E.pyr[l].roi(), E.pyr[l].step(), k.x, k.y, t.umax);
    }
  int total = 0;
  for (int l = 0; l < t.nlevels; ++l) total += (int)E.levelKps[l].size();
  kps.assign(total, KeyPoint());
  desc.assign((size_t)total * 32, 0);
  int mono = 0, stereo = total - 1;
  for (int l = 0; l < t.nlevels; ++l) {
    std::vector<KeyPoint>& lk = E.levelKps[l];
    if (lk.empty()) continue;
    const Level& L = E.pyr[l];
    // clone + GaussianBlur: the clone is continuous, so the border reflects the level itself
    std::vector<u8> src((size_t)L.w * L.h);
    for (int y = 0; y < L.h; ++y) std::memcpy(&src[(size_t)y * L.w], L.roi() + y * L.step(), L.w);
    E.blurred[l].resize(src.size());
    gaussian_blur7_u8(src.data(), L.w, L.h, L.w, E.blurred[l].data(), L.w);
    E.levelDesc[l].resize(lk.size() * 32);
    for (size_t i = 0; i < lk.size(); ++i)
      brief256(E.blurred[l].data(), L.w, lk[i], E.trigMode, &E.levelDesc[l][i * 32]);
    const float s = t.scale[l];
    for (size_t i = 0; i < lk.size(); ++i) {
      KeyPoint k = lk[i];
      if (l != 0) { k.x *= s; k.y *= s; }  // Point2f *= float, :1633
      int dst;
      if (k.x >= lapping[0] && k.x <= lapping[1]) dst = stereo--; else dst = mono++;  // :1637-1646
      kps[dst] = k;
      std::memcpy(&desc[(size_t)dst * 32], &E.levelDesc[l][i * 32], 32);
    }
  }
  if (monoIndex) *monoIndex = mono;
  return total;
}

// =====================================================================================
// Frame glue -- SlamTypes/Frame.cpp:40-206
// =====================================================================================
static const int kGridCols = 64, kGridRows = 48;

struct FrameData {
  int N;
  std::vector<KeyPoint> keys, keysUn;
  std::vector<u8> desc;
  int minX, maxX, minY, maxY;     // image bounds (statics in the reference)
  float gridWInv, gridHInv;
  std::vector<size_t> grid[kGridCols][kGridRows];
};

// ComputeImageBounds, :101-134
static void image_bounds(int W, int H, const float K[4], const float D[4], int b[4]) {
  if (D[0] != 0.0) {
    float c[4][2] = {{0.f, 0.f}, {(float)W, 0.f}, {0.f, (float)H}, {(float)W, (float)H}};
    float u[4][2];
    for (int i = 0; i < 4; ++i) undistort_point(c[i][0], c[i][1], K, D, &u[i][0], &u[i][1]);
    b[0] = (int)std::min(u[0][0], u[2][0]);
    b[1] = (int)std::max(u[1][0], u[3][0]);
    b[2] = (int)std::min(u[0][1], u[1][1]);
    b[3] = (int)std::max(u[2][1], u[3][1]);
  } else {
    b[0] = 0; b[1] = W; b[2] = 0; b[3] = H;
  }
}

static bool pos_in_grid(const FrameData& F, const KeyPoint& kp, int& px, int& py) {  // :89-99
  px = (int)std::round((kp.x - F.minX) * F.gridWInv);
  py = (int)std::round((kp.y - F.minY) * F.gridHInv);
  return !(px < 0 || px >= kGridCols || py < 0 || py >= kGridRows);
}

static void frame_glue(FrameData& F, const std::vector<KeyPoint>& kps, const std::vector<u8>& desc,
                       const float K[4], const float D[4], const int bounds[4]) {
  F.keys = kps; F.desc = desc; F.N = (int)kps.size();
  F.minX = bounds[0]; F.maxX = bounds[1]; F.minY = bounds[2]; F.maxY = bounds[3];
  F.gridWInv = (float)kGridCols / (float)(F.maxX - F.minX);  // :46-47
  F.gridHInv = (float)kGridRows / (float)(F.maxY - F.minY);
  for (int i = 0; i < kGridCols; ++i) for (int j = 0; j < kGridRows; ++j) F.grid[i][j].clear();
  F.keysUn.clear();
  if (kps.empty()) return;  // :61
  F.keysUn = kps;           // UndistortKeyPoints, :136-161
  if (D[0] != 0.0)
    for (int i = 0; i < F.N; ++i) undistort_point(kps[i].x, kps[i].y, K, D, &F.keysUn[i].x, &F.keysUn[i].y);
  for (size_t i = 0; i < F.keysUn.size(); ++i) {  // :71-77
    int gx, gy;
    if (pos_in_grid(F, F.keysUn[i], gx, gy)) F.grid[gx][gy].push_back(i);
  }
}

// GetFeaturesInArea, :163-206
static void features_in_area(const FrameData& F, float x, float y, float r, int minLevel,
                             int maxLevel, std::vector<size_t>& out) {
  out.clear();
  int minCX = std::max(0, (int)std::floor((x - F.minX - r) * F.gridWInv));
  if (minCX >= kGridCols) return;
  int maxCX = std::min(kGridCols - 1, (int)std::ceil((x - F.minX + r) * F.gridWInv));
  if (maxCX < 0) return;
  int minCY = std::max(0, (int)std::floor((y - F.minY - r) * F.gridHInv));
  if (minCY >= kGridRows) return;
  int maxCY = std::min(kGridRows - 1, (int)std::ceil((y - F.minY + r) * F.gridHInv));
  if (maxCY < 0) return;
  const bool check = (minLevel > 0) || (maxLevel >= 0);
  for (int ix = minCX; ix <= maxCX; ++ix)
    for (int iy = minCY; iy <= maxCY; ++iy) {
      const std::vector<size_t>& cell = F.grid[ix][iy];
      for (size_t j = 0; j < cell.size(); ++j) {
        const KeyPoint& k = F.keysUn[cell[j]];
        if (check && !(k.octave >= minLevel && k.octave <= maxLevel)) continue;
        const float dx = k.x - x, dy = k.y - y;
        if (std::fabs(dx) < r && std::fabs(dy) < r) out.push_back(cell[j]);
      }
    }
}

// =====================================================================================
// Matcher -- DBoW2::FORB::distance (Thirdparty/DBoW2/src/FORB.cpp:77-100) and
// ORBmatcher::SearchForInitialization / ComputeThreeMaxima (Features/ORBmatcher.cpp:11-183)
// =====================================================================================
static int hamming256(const u8* a, const u8* b) {
  uint64_t wa[4], wb[4];
  std::memcpy(wa, a, 32); std::memcpy(wb, b, 32);
  uint64_t ret = 0;
  for (int i = 0; i < 4; ++i) {
    uint64_t v = wa[i] ^ wb[i];
    v = v - ((v >> 1) & 0x5555555555555555ULL);
    v = (v & 0x3333333333333333ULL) + ((v >> 2) & 0x3333333333333333ULL);
    v = (v + (v >> 4)) & 0x0F0F0F0F0F0F0F0FULL;
    ret += (v * 0x0101010101010101ULL) >> 56;
  }
  return (int)(double)ret;
}

static const int TH_LOW = 50, HISTO_LENGTH = 30;

static void three_maxima(const std::vector<int>* h, int L, int& i1, int& i2, int& i3) {
  int m1 = 0, m2 = 0, m3 = 0;
  for (int i = 0; i < L; ++i) {
    const int s = (int)h[i].size();
    if (s > m1) { m3 = m2; m2 = m1; m1 = s; i3 = i2; i2 = i1; i1 = i; }
    else if (s > m2) { m3 = m2; m2 = s; i3 = i2; i2 = i; }
    else if (s > m3) { m3 = s; i3 = i; }
  }
  if (m2 < 0.1f * (float)m1) { i2 = -1; i3 = -1; }
  else if (m3 < 0.1f * (float)m1) { i3 = -1; }
}

static int search_for_initialization(const FrameData& F1, const FrameData& F2, float nnratio,
                                     bool checkOri, int windowSize, std::vector<int>& m12,
                                     int counters[3]) {
  int nmatches = 0;
  m12.assign(F1.keysUn.size(), -1);
  std::vector<int> hist[HISTO_LENGTH];
  const float factor = HISTO_LENGTH / 360.0f;
  std::vector<int> matchedDist(F2.N, INT_MAX), m21(F2.N, -1);
  int badDist = 0, badRatio = 0, badOri = 0;
  std::vector<size_t> idx2;
  for (size_t i1 = 0, e = (size_t)F1.N; i1 < e; ++i1) {
    const KeyPoint kp1 = F1.keysUn[i1];
    const int level1 = kp1.octave;
    if (level1 > 0) continue;
    features_in_area(F2, kp1.x, kp1.y, (float)windowSize, level1, level1, idx2);
    if (idx2.empty()) continue;
    const u8* d1 = &F1.desc[i1 * 32];
    int best = INT_MAX, best2 = INT_MAX, bestIdx = -1;
    for (size_t c = 0; c < idx2.size(); ++c) {
      const size_t i2 = idx2[c];
      const int dist = hamming256(d1, &F2.desc[i2 * 32]);
      if (matchedDist[i2] <= dist) continue;
      if (dist < best) { best2 = best; best = dist; bestIdx = (int)i2; }
      else if (dist < best2) best2 = dist;
    }
    if (best > TH_LOW) { ++badDist; continue; }
    if ((float)best > nnratio * (float)best2) { ++badRatio; continue; }
    if (m21[bestIdx] >= 0) { m12[m21[bestIdx]] = -1; --nmatches; }
    m12[i1] = bestIdx; m21[bestIdx] = (int)i1; matchedDist[bestIdx] = best; ++nmatches;
    if (checkOri) {
      float rot = F1.keysUn[i1].angle - F2.keysUn[bestIdx].angle;
      if (rot < 0.0) rot += 360.0f;
      int bin = (int)std::round(rot * factor);
      if (bin == HISTO_LENGTH) bin = 0;
      hist[bin].push_back((int)i1);
    }
  }
  if (checkOri) {
    int a = -1, b = -1, c = -1;
    three_maxima(hist, HISTO_LENGTH, a, b, c);
    for (int i = 0; i < HISTO_LENGTH; ++i)
      if (i != a && i != b && i != c)
        for (size_t k = 0; k < hist[i].size(); ++k) { m12[hist[i][k]] = -1; --nmatches; ++badOri; }
  }
  if (counters) { counters[0] = badDist; counters[1] = badRatio; counters[2] = badOri; }
  return nmatches;
}

}  // namespace orc

// =====================================================================================
// C ABI for ctypes (tests, smoke, bench cpu_baseline)
// =====================================================================================
using namespace orc;

extern "C" {

// ---- primitives (cv2 cross-checks) ----
void orc_resize_linear(const u8* src, int sw, int sh, int sstep, u8* dst, int dw, int dh, int dstep) {
  resize_linear_u8(src, sw, sh, (size_t)sstep, dst, dw, dh, (size_t)dstep);
}
void orc_border101(const u8* src, int w, int h, int sstep, u8* dst, int dstep, int b) {
  copy_make_border101(src, w, h, (size_t)sstep, dst, (size_t)dstep, b);
}
int orc_fast(const u8* img, int cols, int rows, int step, int th, int* xs, int* ys, int* sc, int cap) {
  std::vector<FastPt> v;
  fast9_16_nms(img, cols, rows, (size_t)step, th, v);
  int n = (int)v.size();
  for (int i = 0; i < n && i < cap; ++i) { xs[i] = v[i].x; ys[i] = v[i].y; sc[i] = v[i].score; }
  return n;
}
float orc_fast_atan2(float y, float x) { return fast_atan2(y, x); }
void orc_blur7(const u8* src, int w, int h, int sstep, u8* dst, int dstep) {
  gaussian_blur7_u8(src, w, h, (size_t)sstep, dst, (size_t)dstep);
}
void orc_undistort(const float* pts, int n, const float K[4], const float D[4], float* out) {
  for (int i = 0; i < n; ++i) undistort_point(pts[2 * i], pts[2 * i + 1], K, D, &out[2 * i], &out[2 * i + 1]);
}
void orc_bgr2gray(const u8* bgr, int n, u8* gray) {
  for (int i = 0; i < n; ++i) gray[i] = bgr2gray(bgr[3 * i], bgr[3 * i + 1], bgr[3 * i + 2]);
}
int orc_descriptor_distance(const u8* a, const u8* b) { return hamming256(a, b); }

// ---- extractor ----
void* orc_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
  Extractor* e = new Extractor();
  e->t = make_tables(nfeatures, scaleFactor, nlevels, iniTh, minTh);
  e->trigMode = 2;
  e->policy = 0;
  if (e->t.status) { delete e; return 0; }
  return e;
}
void orc_extractor_destroy(void* h) { delete (Extractor*)h; }
void orc_extractor_set_trig_mode(void* h, int m) { ((Extractor*)h)->trigMode = m; }
// cv::KeyPointsFilter::retainBest as restated in cv_prims.hpp, on n responses: idx_out receives the indices of the kept
// elements in their final order; returns how many
int orc_retain_best(const float* response, int n, int n_points, int* idx_out) {
  struct KP { float response; int idx; };
  std::vector<KP> k((size_t)n);
  for (int i = 0; i < n; ++i) k[i] = KP{response[i], i};
  retain_best(k, n_points);
  for (size_t i = 0; i < k.size(); ++i) idx_out[i] = k[i].idx;
  return (int)k.size();
}
// the restated glibc sinf / cosf (which = 0) or the machine's libm (which = 1) on n floats
void orc_sincosf(int which, const float* y, long long n, float* sn, float* cs) {
  for (long long i = 0; i < n; ++i) {
    if (which == 0) glibc_sincosf_port(y[i], sn + i, cs + i);
    else { sn[i] = sinf(y[i]); cs[i] = cosf(y[i]); }
  }
}
void orc_extractor_set_policy(void* h, int p) { ((Extractor*)h)->policy = p; }
int orc_extractor_tables(void* h, float* scale, float* invScale, float* sigma2, float* invSigma2,
                         int* perLevel, int* umax16) {
  const Tables& t = ((Extractor*)h)->t;
  for (int i = 0; i < t.nlevels; ++i) {
    scale[i] = t.scale[i]; invScale[i] = t.invScale[i]; sigma2[i] = t.sigma2[i];
    invSigma2[i] = t.invSigma2[i]; perLevel[i] = t.perLevel[i];
  }
  for (int i = 0; i < 16; ++i) umax16[i] = t.umax[i];
  return t.nlevels;
}
// returns total keypoints (>=0) or a negative status; kps/desc may be NULL to only run
int orc_extract(void* h, const u8* img, int W, int H, int step, const int lapping[2], KeyPoint* kps,
                u8* desc, int cap, int* monoIndex) {
  Extractor& E = *(Extractor*)h;
  std::vector<KeyPoint> k; std::vector<u8> d;
  int n = extract(E, img, W, H, (size_t)step, lapping, k, d, monoIndex);
  if (n < 0) return n;
  if (kps && desc) {
    int m = std::min(n, cap);
    std::memcpy(kps, k.data(), sizeof(KeyPoint) * (size_t)m);
    std::memcpy(desc, d.data(), (size_t)32 * m);
  }
  return n;
}
int orc_level_size(void* h, int level, int* w, int* hh) {
  Extractor& E = *(Extractor*)h;
  if (level < 0 || level >= (int)E.pyr.size()) return -1;
  *w = E.pyr[level].w; *hh = E.pyr[level].h; return 0;
}
// padded != 0: (w+38)x(h+38) incl. border; else w x h
int orc_get_level(void* h, int level, int padded, u8* out) {
  Extractor& E = *(Extractor*)h;
  const Level& L = E.pyr[level];
  if (padded) { std::memcpy(out, L.pad.data(), L.pad.size()); return 0; }
  for (int y = 0; y < L.h; ++y) std::memcpy(out + (size_t)y * L.w, L.roi() + y * L.step(), L.w);
  return 0;
}
int orc_get_blurred(void* h, int level, u8* out) {
  Extractor& E = *(Extractor*)h;
  if (E.blurred[level].empty()) return 0;
  std::memcpy(out, E.blurred[level].data(), E.blurred[level].size());
  return 1;
}
int orc_get_candidates(void* h, int level, float* xyr, int cap) {
  Extractor& E = *(Extractor*)h;
  int n = (int)E.cands[level].size();
  for (int i = 0; i < n && i < cap; ++i) {
    xyr[3 * i] = E.cands[level][i].x; xyr[3 * i + 1] = E.cands[level][i].y;
    xyr[3 * i + 2] = E.cands[level][i].response;
  }
  return n;
}
int orc_get_level_keypoints(void* h, int level, KeyPoint* kps, u8* desc, int cap) {
  Extractor& E = *(Extractor*)h;
  int n = (int)E.levelKps[level].size();
  int m = std::min(n, cap);
  if (kps) std::memcpy(kps, E.levelKps[level].data(), sizeof(KeyPoint) * (size_t)m);
  if (desc && !E.levelDesc[level].empty()) std::memcpy(desc, E.levelDesc[level].data(), (size_t)32 * m);
  return n;
}
// standalone quadtree (candidates relative to (minX, minY)); returns n kept or negative status
int orc_distribute(const float* xyr, int n, int minX, int maxX, int minY, int maxY, int N, float* out_xyr, int cap) {
  std::vector<Cand> in(n), out;
  for (int i = 0; i < n; ++i) { in[i].x = xyr[3 * i]; in[i].y = xyr[3 * i + 1]; in[i].response = xyr[3 * i + 2]; }
  int rc = distribute_quadtree(in, minX, maxX, minY, maxY, N, out);
  if (rc) return rc;
  int m = (int)out.size();
  for (int i = 0; i < m && i < cap; ++i) { out_xyr[3 * i] = out[i].x; out_xyr[3 * i + 1] = out[i].y; out_xyr[3 * i + 2] = out[i].response; }
  return m;
}

// ---- frame glue + matcher ----
void orc_image_bounds(int W, int H, const float K[4], const float D[4], int bounds[4]) { image_bounds(W, H, K, D, bounds); }

void* orc_frame_create(const KeyPoint* kps, const u8* desc, int n, const float K[4], const float D[4], const int bounds[4]) {
  FrameData* F = new FrameData();
  std::vector<KeyPoint> k(kps, kps + n);
  std::vector<u8> d(desc, desc + (size_t)32 * n);
  frame_glue(*F, k, d, K, D, bounds);
  return F;
}
void orc_frame_destroy(void* f) { delete (FrameData*)f; }
int orc_frame_keys_un(void* f, KeyPoint* out) {
  FrameData& F = *(FrameData*)f;
  std::memcpy(out, F.keysUn.data(), sizeof(KeyPoint) * F.keysUn.size());
  return (int)F.keysUn.size();
}
// CSR grid: offsets[64*48+1] in cell order ix*48+iy, indices ascending per cell
int orc_frame_grid(void* f, unsigned* offsets, unsigned* indices) {
  FrameData& F = *(FrameData*)f;
  unsigned n = 0;
  for (int ix = 0; ix < kGridCols; ++ix)
    for (int iy = 0; iy < kGridRows; ++iy) {
      offsets[ix * kGridRows + iy] = n;
      for (size_t j = 0; j < F.grid[ix][iy].size(); ++j) indices[n++] = (unsigned)F.grid[ix][iy][j];
    }
  offsets[kGridCols * kGridRows] = n;
  return (int)n;
}
int orc_features_in_area(void* f, float x, float y, float r, int minLevel, int maxLevel, int* out, int cap) {
  std::vector<size_t> v;
  features_in_area(*(FrameData*)f, x, y, r, minLevel, maxLevel, v);
  for (size_t i = 0; i < v.size() && (int)i < cap; ++i) out[i] = (int)v[i];
  return (int)v.size();
}
int orc_search_for_initialization(void* f1, void* f2, float nnratio, int checkOri, int windowSize, int* m12, int counters[3]) {
  std::vector<int> m;
  int n = search_for_initialization(*(FrameData*)f1, *(FrameData*)f2, nnratio, checkOri != 0, windowSize, m, counters);
  std::memcpy(m12, m.data(), sizeof(int) * m.size());
  return n;
}

// ---- CPU baseline: frames sharded over host threads; extract every frame, glue, and match
// consecutive pairs (2k, 2k+1).  Returns total matches (a checksum), fills totals. ----
long long orc_run_stream(const u8* imgs, int nframes, int W, int H, int nfeatures, float scaleFactor,
                         int nlevels, int iniTh, int minTh, const float K[4], const float D[4],
                         int windowSize, float nnratio, int threads, long long* totalKps) {
  if (threads < 1) threads = 1;
  int npairs = nframes / 2;
  std::atomic<int> next(0);
  std::atomic<long long> matches(0), kpsum(0);
  int bounds[4];
  image_bounds(W, H, K, D, bounds);
  auto worker = [&]() {
    Extractor E;
    E.t = make_tables(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    E.trigMode = 2;
    E.policy = 0;
    const int lap[2] = {0, 0};
    for (;;) {
      int p = next.fetch_add(1);
      if (p >= npairs) break;
      FrameData F[2];
      for (int s = 0; s < 2; ++s) {
        std::vector<KeyPoint> k; std::vector<u8> d; int mono = 0;
        extract(E, imgs + (size_t)(2 * p + s) * W * H, W, H, (size_t)W, lap, k, d, &mono);
        kpsum += (long long)k.size();
        frame_glue(F[s], k, d, K, D, bounds);
      }
      std::vector<int> m12; int c[3];
      matches += search_for_initialization(F[0], F[1], nnratio, true, windowSize, m12, c);
    }
  };
  std::vector<std::thread> pool;
  for (int i = 0; i < threads; ++i) pool.emplace_back(worker);
  for (size_t i = 0; i < pool.size(); ++i) pool[i].join();
  if (totalKps) *totalKps = kpsum.load();
  return matches.load();
}

// ---- Initializer scoring ---------------------------------------------------
// Initializer::CheckHomography (Initialization/Initializer.cpp:268-353) and CheckFundamental (:355-438) for
// nHyp hypotheses (row-major 3x3 float32 each) on one pair.  The match list mvMatches12 is built as
// Initializer::Initialize does (:27-34): ascending i with matches12[i] >= 0.  All arithmetic is float32 in the
// order written in the reference; `1.0 / x` is a double division narrowed to float (:314,331,293,372).
// Returns N = number of matches.  inliers (may be NULL): nHyp x N bytes.
static int match_list(const KeyPoint* k1, const KeyPoint* k2, const int* m12, int n1, std::vector<float>& uv) {
  uv.clear();
  for (int i = 0; i < n1; ++i)
    if (m12[i] >= 0) {
      uv.push_back(k1[i].x); uv.push_back(k1[i].y); uv.push_back(k2[m12[i]].x); uv.push_back(k2[m12[i]].y);
    }
  return (int)(uv.size() / 4);
}
namespace {
// squared transfer error of point (px, py) mapped by the row-major 3x3 M against (qx, qy); the reciprocal of the
// third coordinate is a double division narrowed to float, as `1.0 / (...)` in the reference (:314, :331)
inline float transfer_err2(const float* M, float px, float py, float qx, float qy) {
  const float wInv = 1.0 / (M[6] * px + M[7] * py + M[8]);
  const float mx = (M[0] * px + M[1] * py + M[2]) * wInv;
  const float my = (M[3] * px + M[4] * py + M[5]) * wInv;
  return (qx - mx) * (qx - mx) + (qy - my) * (qy - my);
}
// squared distance of (qx, qy) to the line (l0, l1, l2)
inline float line_err2(float l0, float l1, float l2, float qx, float qy) {
  const float num = l0 * qx + l1 * qy + l2;
  return num * num / (l0 * l0 + l1 * l1);
}
}  // namespace

int orc_check_homography(const KeyPoint* k1, const KeyPoint* k2, const int* m12, int n1, const float* H21s, const float* H12s,
                         int nHyp, float sigma, float* scores, u8* inliers) {
  std::vector<float> uv;
  const int N = match_list(k1, k2, m12, n1, uv);
  const float chi2Limit = 5.991;                       // :291
  const float invSigma2 = 1.0 / (sigma * sigma);       // :293
  for (int h = 0; h < nHyp; ++h) {
    const float* fwd = H21s + 9 * h;   // frame 1 -> frame 2
    const float* bwd = H12s + 9 * h;   // frame 2 -> frame 1
    float total = 0;
    for (int i = 0; i < N; ++i) {
      const float* q = &uv[4 * i];     // u1, v1, u2, v2
      bool ok = true;
      // frame 2 into frame 1 first (:309-325), then frame 1 into frame 2 (:327-343); two separate additions
      const float c1 = transfer_err2(bwd, q[2], q[3], q[0], q[1]) * invSigma2;
      if (c1 > chi2Limit) ok = false; else total += chi2Limit - c1;
      const float c2 = transfer_err2(fwd, q[0], q[1], q[2], q[3]) * invSigma2;
      if (c2 > chi2Limit) ok = false; else total += chi2Limit - c2;
      if (inliers) inliers[(size_t)h * N + i] = ok ? 1 : 0;
    }
    scores[h] = total;
  }
  return N;
}
int orc_check_fundamental(const KeyPoint* k1, const KeyPoint* k2, const int* m12, int n1, const float* F21s, int nHyp, float sigma,
                          float* scores, u8* inliers) {
  std::vector<float> uv;
  const int N = match_list(k1, k2, m12, n1, uv);
  const float chi2Limit = 3.841;   // :372, one degree of freedom for the test
  const float chi2Score = 5.991;   // :374, the score keeps the homography's constant
  const float invSigma2 = 1.0 / (sigma * sigma);
  for (int h = 0; h < nHyp; ++h) {
    const float* F = F21s + 9 * h;
    float total = 0;
    for (int i = 0; i < N; ++i) {
      const float* q = &uv[4 * i];
      bool ok = true;
      // epipolar line of point 1 in image 2: F x1 (:392-404)
      const float c1 = line_err2(F[0] * q[0] + F[1] * q[1] + F[2], F[3] * q[0] + F[4] * q[1] + F[5],
                                 F[6] * q[0] + F[7] * q[1] + F[8], q[2], q[3]) * invSigma2;
      if (c1 > chi2Limit) ok = false; else total += chi2Score - c1;
      // epipolar line of point 2 in image 1: F^T x2 (:414-426)
      const float c2 = line_err2(F[0] * q[2] + F[3] * q[3] + F[6], F[1] * q[2] + F[4] * q[3] + F[7],
                                 F[2] * q[2] + F[5] * q[3] + F[8], q[0], q[1]) * invSigma2;
      if (c2 > chi2Limit) ok = false; else total += chi2Score - c2;
      if (inliers) inliers[(size_t)h * N + i] = ok ? 1 : 0;
    }
    scores[h] = total;
  }
  return N;
}
// the selection of the RANSAC loops FindHomography / FindFundamental (:205-209, :257-261): the running best starts
// at 0 and is replaced only by a strictly larger score; -1 when no hypothesis scores above 0
int orc_ransac_best(const float* scores, int nHyp) {
  float score = 0; int best = -1;
  for (int it = 0; it < nHyp; ++it)
    if (scores[it] > score) { score = scores[it]; best = it; }
  return best;
}

}  // extern "C"
