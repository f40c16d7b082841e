// oracle/orb_port_impl.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see orb_port.hpp).
#pragma once
#include <list>

namespace orbport {

namespace detail {
struct Node {
  int x0, x1, y0, y1;     // UL.x, UR.x, UL.y, BL.y (ORBextractor.h:41)
  std::vector<int> keys;  // indices into the candidate array, in candidate order
  bool frozen = false;    // bNoMore
  long seq = 0;           // creation sequence == address order under a monotonic allocator
};
typedef std::list<Node>::iterator NodeIt;

// ExtractorNode::DivideNode, ORBextractor.cc:481-537
inline void split(const Node& p, const std::vector<Candidate>& c, Node ch[4]) {
  const int halfX = (int)ceil((float)(p.x1 - p.x0) / 2);
  const int halfY = (int)ceil((float)(p.y1 - p.y0) / 2);
  const int mx = p.x0 + halfX, my = p.y0 + halfY;
  ch[0] = Node{p.x0, mx, p.y0, my, {}, false, 0};
  ch[1] = Node{mx, p.x1, p.y0, my, {}, false, 0};
  ch[2] = Node{p.x0, mx, my, p.y1, {}, false, 0};
  ch[3] = Node{mx, p.x1, my, p.y1, {}, false, 0};
  for (int k : p.keys) {
    const bool right = !((float)c[k].x < (float)mx), low = !((float)c[k].y < (float)my);
    ch[(low ? 2 : 0) + (right ? 1 : 0)].keys.push_back(k);
  }
  for (int q = 0; q < 4; ++q) ch[q].frozen = ch[q].keys.size() == 1;
}
}  // namespace detail

// DistributeOctTree, ORBextractor.cc:539-763.  boxW/boxH = maxX-minX / maxY-minY.
inline void Extractor::distribute(const std::vector<Candidate>& keys, int boxW, int boxH, int N,
                                  std::vector<Candidate>& kept) const {
  using namespace detail;
  kept.clear();
  const int nIni = (int)round((float)boxW / (float)boxH);
  const float hX = (float)boxW / nIni;
  std::list<Node> nodes;
  long seq = 0;
  std::vector<NodeIt> roots;
  for (int i = 0; i < nIni; ++i) {
    Node n{(int)(hX * (float)i), (int)(hX * (float)(i + 1)), 0, boxH, {}, false, seq++};
    nodes.push_back(n);
    roots.push_back(std::prev(nodes.end()));
  }
  for (int k = 0; k < (int)keys.size(); ++k) roots[(size_t)((float)keys[k].x / hX)]->keys.push_back(k);
  for (NodeIt it = nodes.begin(); it != nodes.end();) {
    if (it->keys.size() == 1) { it->frozen = true; ++it; }
    else if (it->keys.empty()) it = nodes.erase(it);
    else ++it;
  }

  typedef std::pair<int, NodeIt> Pending;  // (key count, node); node order == seq order
  std::vector<Pending> pending;
  auto add_children = [&](Node ch[4]) {
    for (int q = 0; q < 4; ++q) {
      if (ch[q].keys.empty()) continue;
      ch[q].seq = seq++;
      nodes.push_front(ch[q]);
      if (ch[q].keys.size() > 1) pending.push_back(Pending((int)ch[q].keys.size(), nodes.begin()));
    }
  };

  bool finish = false;
  while (!finish) {
    int prevSize = (int)nodes.size();
    int nToExpand = 0;
    pending.clear();
    for (NodeIt it = nodes.begin(); it != nodes.end();) {
      if (it->frozen) { ++it; continue; }
      Node ch[4];
      split(*it, keys, ch);
      size_t before = pending.size();
      add_children(ch);
      nToExpand += (int)(pending.size() - before);
      it = nodes.erase(it);
    }
    if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
      finish = true;
    } else if ((int)nodes.size() + nToExpand * 3 > N) {
      while (!finish) {
        prevSize = (int)nodes.size();
        std::vector<Pending> prev = pending;
        pending.clear();
        std::sort(prev.begin(), prev.end(), [](const Pending& a, const Pending& b) {
          return a.first != b.first ? a.first < b.first : a.second->seq < b.second->seq;
        });
        for (int j = (int)prev.size() - 1; j >= 0; --j) {
          Node ch[4];
          split(*prev[j].second, keys, ch);
          add_children(ch);
          nodes.erase(prev[j].second);
          if ((int)nodes.size() >= N) break;
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finish = true;
      }
    }
  }
  // best response per node, first wins ties (ORBextractor.cc:744-760)
  for (const Node& n : nodes) {
    int best = n.keys[0];
    for (size_t k = 1; k < n.keys.size(); ++k)
      if (keys[n.keys[k]].score > keys[best].score) best = n.keys[k];
    kept.push_back(keys[best]);
  }
}

// IC_Angle, ORBextractor.cc:77-104 (x,y are level-image coordinates)
inline float Extractor::ic_angle(int l, int x, int y) const {
  const Level& L = pyr[l];
  const ptrdiff_t step = (ptrdiff_t)L.step;
  const uint8_t* c = L.roi() + (ptrdiff_t)y * step + x;
  int m01 = 0, m10 = 0;
  for (int u = -HALF_PATCH; u <= HALF_PATCH; ++u) m10 += u * c[u];
  for (int v = 1; v <= HALF_PATCH; ++v) {
    int vsum = 0, d = umax[v];
    for (int u = -d; u <= d; ++u) {
      int p = c[u + v * step], m = c[u - v * step];
      vsum += p - m;
      m10 += u * (p + m);
    }
    m01 += v * vsum;
  }
  return cvprims::fast_atan2_deg((float)m01, (float)m10);
}

// computeOrbDescriptor, ORBextractor.cc:107-147.  `blurred` points at pixel (0,0) of the
// blurred level image.
inline void Extractor::describe(const uint8_t* blurred, size_t step, int x, int y, float angle_deg,
                                uint8_t* desc) const {
  const float factorPI = (float)(3.14159265358979323846 / 180.f);
  float angle = angle_deg * factorPI;
  float a = (float)cos(angle), b = (float)sin(angle);  // double cos/sin of a float, rounded
  const uint8_t* c = blurred + (ptrdiff_t)y * (ptrdiff_t)step + x;
  auto sample = [&](int idx) -> int {
    const float px = (float)kPattern[2 * idx], py = (float)kPattern[2 * idx + 1];
    int yy = cvprims::cv_round(px * b + py * a);
    int xx = cvprims::cv_round(px * a - py * b);
    return c[(ptrdiff_t)yy * (ptrdiff_t)step + xx];
  };
  for (int i = 0; i < 32; ++i) {
    int val = 0;
    for (int k = 0; k < 8; ++k) val |= (sample(16 * i + 2 * k) < sample(16 * i + 2 * k + 1)) << k;
    desc[i] = (uint8_t)val;
  }
}

// operator(), ORBextractor.cc:1051-1113
inline void Extractor::extract(const uint8_t* img, int w, int h, size_t stride,
                               std::vector<KeyPoint>& kps, std::vector<uint8_t>& desc) {
  kps.clear();
  desc.clear();
  if (!img || w <= 0 || h <= 0) return;
  compute_pyramid(img, w, h, stride);
  lastCandidates.assign(nlevels, {});
  lastKept.assign(nlevels, {});
  std::vector<uint8_t> blur;
  for (int l = 0; l < nlevels; ++l) {
    const CellGrid g = cell_grid(l);
    detect_level(l, lastCandidates[l]);
    distribute(lastCandidates[l], g.maxBX - g.minB, g.maxBY - g.minB, quota[l], lastKept[l]);
  }
  for (int l = 0; l < nlevels; ++l) {
    const std::vector<Candidate>& kept = lastKept[l];
    if (kept.empty()) continue;
    const Level& L = pyr[l];
    const CellGrid g = cell_grid(l);
    blur.assign((size_t)L.w * L.h, 0);
    cvprims::gaussian_blur7_u8(L.roi(), L.w, L.h, L.step, blur.data(), (size_t)L.w);
    const int patch = (int)(PATCH_SIZE * scale[l]);
    for (const Candidate& c : kept) {
      KeyPoint k;
      const int x = c.x + g.minB, y = c.y + g.minB;
      k.x = (float)x; k.y = (float)y;
      k.size = (float)patch;
      k.angle = ic_angle(l, x, y);
      k.response = (float)c.score;
      k.octave = l;
      k.class_id = -1;
      size_t o = desc.size();
      desc.resize(o + 32);
      describe(blur.data(), (size_t)L.w, x, y, k.angle, &desc[o]);
      if (l != 0) { k.x *= scale[l]; k.y *= scale[l]; }
      kps.push_back(k);
    }
  }
}

}  // namespace orbport
