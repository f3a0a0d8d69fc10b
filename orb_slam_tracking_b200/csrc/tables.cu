// Host tables of a context: every float -> int decision the reference makes on the CPU (level
// sizes, resize coefficients, per-level quotas, umax, cell geometry, quadtree roots) is made here
// once, on the host, with the reference's own expressions, and uploaded as integers.  The GPU
// never re-derives them.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "ctx.cuh"

namespace orbx {

static inline int cv_round_f(float v) { return (int)lrintf(v); }    // cvRound: half-to-even
static inline int cv_round_d(double v) { return (int)lrint(v); }
static inline int cv_floor_f(float v) { int i = (int)v; return i - (i > v); }
static inline int cv_ceil_f(float v) { int i = (int)v; return i + (i < v); }

// ORBextractor::ORBextractor, Features/ORBextractor.cpp:492-595.  Returns ORBX_ERR_CTOR where the
// reference calls exit(1).
int compute_tables(int nfeatures, float scaleFactorF, int nlevels, HostTables& t) {
  const double scaleFactor = scaleFactorF;  // the member is a double holding the float (hpp:142)
  if (nlevels < 1 || nlevels > kMaxLevels) return ORBX_ERR_ARG;
  if (scaleFactor == 1.0 && nlevels > 1) return ORBX_ERR_CTOR;  // :502-505
  t.scale.assign(nlevels, 1.0f);
  t.sigma2.assign(nlevels, 1.0f);
  for (int i = 1; i < nlevels; ++i) {
    t.scale[i] = (float)(t.scale[i - 1] * scaleFactor);  // :512
    t.sigma2[i] = t.scale[i] * t.scale[i];                // :513
  }
  t.invScale.resize(nlevels);
  t.invSigma2.resize(nlevels);
  for (int i = 0; i < nlevels; ++i) {
    t.invScale[i] = 1.0f / t.scale[i];
    t.invSigma2[i] = 1.0f / t.sigma2[i];
  }
  // quotas: geometric series in float, cvRound per level, remainder to the last level (:529-548)
  t.perLevel.assign(nlevels, 0);
  const float factor = (float)(1.0f / scaleFactor);
  float want = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)(float)nlevels));
  int sum = 0;
  for (int l = 0; l < nlevels - 1; ++l) {
    t.perLevel[l] = cv_round_f(want);
    sum += t.perLevel[l];
    want *= factor;
  }
  t.perLevel[nlevels - 1] = nfeatures - sum > 0 ? nfeatures - sum : 0;
  // umax (:562-594)
  std::memset(t.umax, 0, sizeof(t.umax));
  const int vmax = cv_floor_f(kHalfPatch * std::sqrt(2.f) / 2 + 1);
  const int vmin = cv_ceil_f(kHalfPatch * std::sqrt(2.f) / 2);
  const double hp2 = kHalfPatch * kHalfPatch;
  for (int v = 0; v <= vmax; ++v) t.umax[v] = cv_round_d(std::sqrt(hp2 - v * v));
  for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
    while (t.umax[v0] == t.umax[v0 + 1]) ++v0;
    t.umax[v] = v0;
    ++v0;
  }
  return ORBX_OK;
}

// cv::resize INTER_LINEAR 8U coefficient tables.  Layout: xofs[dw] | xcoef[dw] | { r0 | r1 << 16, ycoef }[dh]:
// the two source rows of an output row already clipped to [0, sh - 1] like OpenCV's row loop does at use.
void build_resize_tab(int sw, int sh, int dw, int dh, std::vector<int>& tab) {
  tab.assign(2 * (size_t)dw + 2 * (size_t)dh, 0);
  const double sxs = 1. / ((double)dw / sw), sys = 1. / ((double)dh / sh);
  for (int d = 0; d < dw; ++d) {
    float f = (float)((d + 0.5) * sxs - 0.5);
    int s = cv_floor_f(f);
    f -= s;
    if (s < 0) { f = 0; s = 0; }
    if (s >= sw - 1) { f = 0; s = sw - 1; }
    const int a0 = cv_round_f((1.f - f) * 2048.f), a1 = cv_round_f(f * 2048.f);
    tab[d] = s;
    tab[dw + d] = (a0 & 0xffff) | (a1 << 16);
  }
  for (int d = 0; d < dh; ++d) {
    float f = (float)((d + 0.5) * sys - 0.5);
    const int s = cv_floor_f(f);
    f -= s;
    const int b0 = cv_round_f((1.f - f) * 2048.f), b1 = cv_round_f(f * 2048.f);
    const int r0 = std::min(std::max(s, 0), sh - 1), r1 = std::min(std::max(s + 1, 0), sh - 1);
    tab[2 * dw + 2 * d] = r0 | (r1 << 16);
    tab[2 * dw + 2 * d + 1] = (b0 & 0xffff) | (b1 << 16);
  }
}

static inline u32 align_up(u32 v, u32 a) { return (v + a - 1) / a * a; }

// Fills P.lv[], the cell table, the blur tile table and the quadtree root tables.
// Returns ORBX_ERR_ARG for shapes on which the reference itself is undefined (a level narrower
// than 67 px gives nCols == 0 and a division by zero at :1073) and ORBX_ERR_ASPECT where
// DistributeOctTree would index an empty root vector (:706, aspect ratio rounding to 0).
int build_geometry(Params& P, const HostTables& T, int fastTileBudget, std::vector<StripDesc>& strips, std::vector<BlurTile>& tiles,
                   std::vector<u8>& rootTab, std::vector<int>& rootEdges) {
  u32 imgOff = 0, candOff = 0, nodeOff = 0;
  int kpOff = 0;
  strips.clear(); tiles.clear(); rootTab.clear(); rootEdges.clear();
  int tileMax = 0, listW = 0, listP = 0;
  for (int l = 0; l < P.nlevels; ++l) {
    LevelInfo& L = P.lv[l];
    L.w = cv_round_f((float)P.W * T.invScale[l]);  // :1662-1663
    L.h = cv_round_f((float)P.H * T.invScale[l]);
    L.pitch = (int)align_up((u32)L.w, 128);
    L.imgOff = imgOff;
    imgOff += align_up((u32)L.pitch * (u32)L.h, 256);
    L.scale = T.scale[l];
    L.patchSize = (float)(int)(kPatch * T.scale[l]);  // :1165
    L.quota = T.perLevel[l];
    L.kpOff = kpOff;
    kpOff += L.quota;
    // cell grid, :1056-1075
    const int minBX = kMinBorder, minBY = kMinBorder;
    const int maxBX = L.w - kEdge + 3, maxBY = L.h - kEdge + 3;
    const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
    L.distW = maxBX - minBX;
    L.distH = maxBY - minBY;
    L.nCols = (int)(width / 35.f);
    L.nRows = (int)(height / 35.f);
    if (L.nCols < 1 || L.nRows < 1 || L.w > 32000 || L.h > 32000) return ORBX_ERR_ARG;
    L.wCell = (int)std::ceil(width / L.nCols);
    L.hCell = (int)std::ceil(height / L.nRows);
    if (L.wCell + 6 > kMaxCellDim || L.hCell + 6 > kMaxCellDim) return ORBX_ERR_ARG;
    const u32 cellMul = ((1u << kCellShift) - 1) / (u32)L.wCell + 1;  // exact floor(xb / wCell) for xb * wCell < 2^22 (strip interiors are < 1024 wide)
    // The cells of a cell row that cv::FAST can find anything in are j = 0 .. nValid - 1 (the skipped and the
    // degenerate ones are at the end, :1101); their tested interiors tile [19, maxBX - 3) without gaps.
    int nValid = 0, lastMaxX = 0;
    for (int j = 0; j < L.nCols; ++j) {
      const float iniX = (float)(minBX + j * L.wCell);
      float maxX = iniX + L.wCell + 6;
      if (iniX >= maxBX - 6) continue;  // :1101
      if (maxX > maxBX) maxX = (float)maxBX;
      if ((int)maxX - (int)iniX - 6 < 1) continue;  // cv::FAST yields nothing for ROIs under 7 px
      if (j != nValid) return ORBX_ERR_ARG;  // cannot happen: skipped cells only follow valid ones
      nValid = j + 1;
      lastMaxX = (int)maxX;
    }
    // k_fast works on strips: a cell row cut into nSplit runs of cells whose staged tile fits the shared-memory budget.
    // Staged row: up to 16 bytes of alignment apron + ROI + the word right of the last interior word, as an odd multiple
    // of 16 bytes (rows 8 apart then share a bank instead of every row).
    const int boxH = std::min(L.hCell + 6, maxBY - minBY);
    int nSplit = 1, perStrip = nValid;
    auto box_w = [&](int nc) {
      u32 w = align_up((u32)(nc * L.wCell + 6 + 20), 16);
      if (((w / 16) & 1) == 0) w += 16;
      return (int)w;
    };
    for (;; ++nSplit) {
      perStrip = (nValid + nSplit - 1) / nSplit;
      // phase A of k_fast gives one lane to each aligned pair of interior words: at most 32 pairs (244 + 6 columns) per strip
      if (perStrip <= 1 || (perStrip <= 24 && perStrip * L.wCell <= 244 && box_w(perStrip) * boxH <= fastTileBudget)) break;
    }
    if (nValid < 1) return ORBX_ERR_ARG;
    L.fastBoxW = box_w(perStrip);
    L.fastBoxH = boxH;
    if (L.fastBoxW > 508 || boxH > 127 || perStrip * L.wCell > 244) return ORBX_ERR_ARG;  // list entries: 9 bits of byte column, 7 bits of row
    tileMax = std::max(tileMax, L.fastBoxW * L.fastBoxH);
    int cap = 0;
    for (int i = 0; i < L.nRows; ++i) {
      const float iniY = (float)(minBY + i * L.hCell);
      float maxY = iniY + L.hCell + 6;
      if (iniY >= maxBY - 3) continue;  // :1088
      if (maxY > maxBY) maxY = (float)maxBY;
      const int ih = (int)maxY - (int)iniY - 6;
      if (ih < 1) continue;
      for (int j0 = 0; j0 < nValid; j0 += perStrip) {
        const int j1 = std::min(nValid, j0 + perStrip);
        StripDesc s;
        const int x0 = minBX + j0 * L.wCell;
        const int x1 = j1 == nValid ? lastMaxX : minBX + (j1 - 1) * L.wCell + L.wCell + 6;
        // the tile starts at the 16-byte aligned column xs <= x0 - 1 of the level, so byte b of a staged row <-> level
        // column xs + b, and the word left of the first interior word is always staged
        const int xs = (x0 - 1) & ~15, off = x0 - xs;
        s.level = (short)l; s.nCells = (short)(j1 - j0);
        s.tmaX = (short)(xs >> 2); s.y0 = (short)iniY;
        s.SB = (short)L.fastBoxW;
        s.bLo = (short)(off + 3); s.bHi = (short)(off + (x1 - x0) - 3);
        s.nR = (short)ih;
        s.wCell = (short)L.wCell;
        s.sx = (short)(j0 * L.wCell); s.sy = (short)(i * L.hCell);
        s.nIniM1 = 0;  // level fields that are not known yet are filled in below
        s.tileBytes = (L.fastBoxW * L.fastBoxH) | (0 << 24) | (7 << 28);
        s.cellMul = cellMul;
        s.index0 = i * L.nCols + j0;
        s.candOff = 0; s.candCap = 0; s.rootTabOff = 0;
        strips.push_back(s);
        // per-warp list segments of k_fast (rows r = warp, warp + kFastWarps, ...): every word / pixel of its rows may be listed
        const int iwS = x1 - x0 - 6, rowsPerWarp = (ih + kFastWarps - 1) / kFastWarps;
        listW = std::max(listW, rowsPerWarp * (iwS / 4 + 4));
        listP = std::max(listP, rowsPerWarp * iwS);
        for (int j = j0; j < j1; ++j) {
          const int iw = std::min(L.wCell, x1 - 3 - (x0 + 3 + (j - j0) * L.wCell));
          cap += ((iw + 1) / 2) * ((ih + 1) / 2);  // NMS survivors are never 8-adjacent
        }
      }
    }
    if ((u32)(L.nCols * L.nRows) >= (1u << 18)) return ORBX_ERR_ARG;  // order key has 18 cell bits
    L.candCap = cap + cap / 8 + 64;  // the bound of this grid, with room for the coarser grid of ComputeKeyPointsOld
    L.candOff = candOff;
    candOff += align_up((u32)L.candCap, 32);
    // quadtree roots, :706-739, and the key -> root map, :743-748
    L.nIni = (int)std::round((float)(maxBX - minBX) / (maxBY - minBY));
    if (L.nIni < 1) return ORBX_ERR_ASPECT;
    if (L.nIni > 255) return ORBX_ERR_ARG;
    const float hX = (float)(maxBX - minBX) / L.nIni;
    L.rootEdgeOff = (u32)rootEdges.size();
    for (int i = 0; i <= L.nIni; ++i) rootEdges.push_back((int)(hX * (float)i));
    L.rootTabOff = (u32)rootTab.size();
    for (int x = 0; x <= L.distW; ++x) {
      int r = (int)((float)x / hX);
      if (r >= L.nIni) r = L.nIni - 1;  // unreachable for real candidates (x < distW - 3)
      rootTab.push_back((u8)r);
    }
    L.nodeCap = 8 * L.quota + 64 + 4 * L.nIni;
    if (L.nodeCap >= (1 << 24) || L.candCap >= (1 << 24)) return ORBX_ERR_ARG;  // packed sort words: 24-bit count and node id
    L.nodeOff = nodeOff;
    for (StripDesc& sd : strips)
      if (sd.level == l) { sd.nIniM1 = (short)(L.nIni - 1); sd.candOff = L.candOff; sd.candCap = L.candCap; sd.rootTabOff = L.rootTabOff; }
    nodeOff += align_up((u32)L.nodeCap, 32);
    // blur strips of 120 x 64 (one warp each)
    for (int y = 0; y < L.h; y += kBlurTH)
      for (int x = 0; x < L.w; x += kBlurTW) {
        BlurTile t;
        t.level = (short)l; t.pad = 0; t.x0 = (short)x; t.y0 = (short)y;
        tiles.push_back(t);
      }
  }
  P.slabBytes = imgOff;
  P.candTotal = candOff;
  P.nodeTotal = nodeOff;
  P.nStrips = (int)strips.size();
  P.fastTileMax = (int)align_up((u32)tileMax, 128);
  P.fastSegW = (listW + 7) & ~7;
  P.fastSegP = (listP + 7) & ~7;
  P.nBlurTiles = (int)tiles.size();
  return ORBX_OK;
}

// Strip table of k_fast for ComputeKeyPointsOld's cell grid (Features/ORBextractor.cpp:1222-1330): levelCols x levelRows
// cells of cellW x cellH tested pixels starting at (19, 19); the last column / row runs to the border; a cell is skipped
// when nothing is left for it (hX <= 0, :1317-1320).  Same strip rules as above (whole cells, at most 244 tested columns).
// Fills the o* fields of P.lv[]; returns ORBX_ERR_ARG where the reference itself divides by zero (levelCols == 0) or a
// cell does not fit one strip.
int build_strips_old(Params& P, const HostTables& T, int fastTileBudget, std::vector<StripDesc>& strips, int& tileMaxOut, int& segWOut,
                     int& segPOut, int& cellTotal) {
  strips.clear();
  int tileMax = 0, listW = 0, listP = 0;
  cellTotal = 0;
  const float imageRatio = (float)P.lv[0].w / (float)P.lv[0].h;  // :1213-1214
  for (int l = 0; l < P.nlevels; ++l) {
    LevelInfo& L = P.lv[l];
    const int nDesired = T.perLevel[l];
    L.oCols = (int)std::sqrt((float)nDesired / (5 * imageRatio));
    L.oRows = (int)(imageRatio * L.oCols);
    if (L.oCols < 1 || L.oRows < 1 || L.oCols * L.oRows > 4096) return ORBX_ERR_ARG;
    const int minB = kEdge, maxBX = L.w - kEdge, maxBY = L.h - kEdge;
    const int W = maxBX - minB, H = maxBY - minB;
    L.oCellW = (int)std::ceil((float)W / L.oCols);
    L.oCellH = (int)std::ceil((float)H / L.oRows);
    L.oNfCell = (int)std::ceil((float)nDesired / (L.oCols * L.oRows));
    if (L.oCellW > 480 || L.oCellH + 6 > 127) return ORBX_ERR_ARG;  // a cell is one strip at least: 9-bit byte columns, 7-bit rows
    // A cell row / column other than the last ends at minB + (i + 1) * cell + 3 (:1279, :1309).  Beyond maxB + 3 the reference tests and
    // keeps keypoints closer than 19 px to the level's border: IC_Angle then needs the 19-px frame (not stored here), the descriptor
    // reads outside the blurred level (undefined in the reference), and a cell beyond the image makes cv::Mat::rowRange assert.
    // Those shapes (very wide or very tall levels with many small cells) are outside the function's working range and are refused.
    if ((L.oRows - 1) * L.oCellH > H || (L.oCols - 1) * L.oCellW > W) return ORBX_ERR_ARG;
    cellTotal += L.oCols * L.oRows;
    // columns with something to test: ROI [16 + j cellW, 16 + j cellW + hX), hX = cellW + 6 or (last column) w - 16 - iniX
    int nValid = 0, lastX1 = 0;
    for (int j = 0; j < L.oCols; ++j) {
      const int iniX = minB + j * L.oCellW - 3;
      const int hX = j == L.oCols - 1 ? maxBX + 3 - iniX : L.oCellW + 6;
      if (hX <= 0) continue;  // only the last column can be skipped
      if (hX - 6 < 1) continue;  // cv::FAST yields nothing for ROIs under 7 px (the cell stays in the books with 0 keypoints)
      nValid = j + 1;
      lastX1 = iniX + hX;
    }
    // (a 16-bit reciprocal is exact only for xb * cellW < 65536: the last column of a 343-px cell of ComputeKeyPointsOld landed in
    // the next cell -- found by tools/gpu/fuzz_stream.py)
    const u32 cellMul = ((1u << kCellShift) - 1) / (u32)L.oCellW + 1;
    int nSplit = 1, perStrip = nValid;
    auto box_w = [&](int nc) {
      u32 w = align_up((u32)(nc * L.oCellW + 6 + 20), 16);
      if (((w / 16) & 1) == 0) w += 16;
      return (int)w;
    };
    const int boxH = L.oCellH + 6;
    for (; nValid > 0; ++nSplit) {
      perStrip = (nValid + nSplit - 1) / nSplit;
      if (perStrip <= 1 || (perStrip <= 24 && perStrip * L.oCellW <= 244 && box_w(perStrip) * boxH <= fastTileBudget)) break;
    }
    if (nValid > 0 && perStrip * L.oCellW > 480) return ORBX_ERR_ARG;
    L.oBoxW = box_w(std::max(perStrip, 1));
    L.oBoxH = boxH;
    if (L.oBoxW > 508) return ORBX_ERR_ARG;
    tileMax = std::max(tileMax, L.oBoxW * L.oBoxH);
    for (int i = 0; i < L.oRows && nValid > 0; ++i) {
      const int iniY = minB + i * L.oCellH - 3;
      const int hY = i == L.oRows - 1 ? maxBY + 3 - iniY : L.oCellH + 6;
      if (hY <= 0) continue;
      const int ih = hY - 6;
      if (ih < 1) continue;
      for (int j0 = 0; j0 < nValid; j0 += perStrip) {
        const int j1 = std::min(nValid, j0 + perStrip);
        StripDesc s;
        const int x0 = minB - 3 + j0 * L.oCellW;
        const int x1 = j1 == nValid ? lastX1 : minB - 3 + j1 * L.oCellW + 6;
        const int xs = (x0 - 1) & ~15, off = x0 - xs;
        s.level = (short)l; s.nCells = (short)(j1 - j0);
        s.tmaX = (short)(xs >> 2); s.y0 = (short)iniY;
        s.SB = (short)L.oBoxW;
        s.bLo = (short)(off + 3); s.bHi = (short)(off + (x1 - x0) - 3);
        s.nR = (short)ih;
        s.wCell = (short)L.oCellW;
        s.sx = (short)(j0 * L.oCellW); s.sy = (short)(i * L.oCellH);
        s.nIniM1 = (short)(L.nIni - 1);
        s.tileBytes = (L.oBoxW * L.oBoxH) | (3 << 24) | (9 << 28);
        s.cellMul = cellMul;
        s.index0 = i * L.oCols + j0;
        s.candOff = L.candOff; s.candCap = L.candCap; s.rootTabOff = L.rootTabOff;
        strips.push_back(s);
        const int iwS = x1 - x0 - 6, rowsPerWarp = (ih + kFastWarps - 1) / kFastWarps;
        listW = std::max(listW, rowsPerWarp * (iwS / 4 + 4));
        listP = std::max(listP, rowsPerWarp * iwS);
      }
    }
  }
  tileMaxOut = (int)align_up((u32)tileMax, 128);
  segWOut = (listW + 7) & ~7;
  segPOut = (listP + 7) & ~7;
  return ORBX_OK;
}

}  // namespace orbx
