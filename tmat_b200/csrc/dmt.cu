// Discrete-Morse-theory graph of a grayscale image.
// Reference: fl_tissue_model_tools/dmtgraph.py:38-453 (compute_dmt_graph; its native twin is
// pydmtgraph/pydmtgraph/dmtgraph/DMTGraph.cpp:232-615), called at scripts/compute_branches.py:424 on the <= 384 px
// centerline-weighted prediction.  The algorithm is two union-find sweeps over the edges of the triangulated pixel grid
// in filtration order plus two breadth-first passes: sequential by construction (every merge depends on all earlier
// ones), so this is NATIVE HOST code -- C++ inside libtmat_b200, no CUDA kernel and nothing for a GPU to do at this size
// (440 k edges at 384^2; the reference's numba version takes seconds, this one tens of milliseconds).  Every tie-break,
// float32 comparison and quirk of the reference is kept so that vertices and edges come out in the same order:
//   * the image is negated; vertices with |value| <= float32(1e-8) (np.isclose(v, 0)) drop all their edges (:73-79);
//   * edges are created vertical, horizontal, diagonal (:199-276) and sorted by (max endpoint value, creation index);
//     the second sweep runs over exactly the reverse order (lexsort of the negated keys, :92-95);
//   * merge() keeps the root with the smaller (value, index) -- larger for the dual sweep -- (:124-152);
//   * a vertex remembers at most four tree neighbours; a fifth is dropped silently (:338-346);
//   * vertex and edge output order is the order of first appearance while walking the sorted edges (:431-451).
#include "common.cuh"
#include <algorithm>
#include <cmath>
#include <limits>
#include <numeric>

namespace {

struct DmtEdge {
  int v1, v2, dv1, dv2;
  int pair_type;        // 0 unknown, 1 vertex-edge, 2 edge-triangle
  float persistence;
  float max_val;
};

int uf_find_host(std::vector<int>& parent, int v) {
  int root = v;
  while (parent[root] != root) root = parent[root];
  while (parent[v] != v) { int p = parent[v]; parent[v] = root; v = p; }     // compression does not change any result
  return root;
}
// returns NaN when both ends are already in one component, else the value of the root that stops being a root
float uf_merge_host(std::vector<int>& parent, const std::vector<float>& val, int a, int b, bool flip) {
  const int p1 = uf_find_host(parent, a), p2 = uf_find_host(parent, b);
  if (p1 == p2) return std::numeric_limits<float>::quiet_NaN();
  const bool first_wins = flip ? (val[p1] > val[p2] || (val[p1] == val[p2] && p1 > p2))
                               : (val[p1] < val[p2] || (val[p1] == val[p2] && p1 < p2));
  if (first_wins) { parent[p2] = p1; return val[p2]; }
  parent[p1] = p2;
  return val[p1];
}

}  // namespace

extern "C" int tmat_dmt_graph_host(const float* img_host, int H, int W, float delta1, float delta2,
                                   int32_t* vertices_host, int64_t cap_vertices, int64_t* n_vertices,
                                   int32_t* edges_host, int64_t cap_edges, int64_t* n_edges_out) {
  TMAT_REQUIRE(img_host && n_vertices && n_edges_out, "tmat_dmt_graph: NULL argument");
  TMAT_REQUIRE(H >= 2 && W >= 2, "tmat_dmt_graph: image %dx%d too small", H, W);
  TMAT_REQUIRE((int64_t)H * W < (1 << 24), "tmat_dmt_graph: the reference stores indices in float32; %dx%d exceeds 2^24 vertices", H, W);
  const int n_verts = H * W;
  const int n_dual = (H - 1) * (W - 1) * 2;               // + 1 outer face at index n_dual
  std::vector<float> val(n_verts);
  for (int i = 0; i < n_verts; ++i) val[i] = -img_host[i];
  std::vector<float> dval(n_dual + 1);
  for (int r = 0; r + 1 < H; ++r)
    for (int c = 0; c + 1 < W; ++c) {
      const float a = val[r * W + c], b = val[r * W + c + 1], d = val[(r + 1) * W + c], e = val[(r + 1) * W + c + 1];
      // np.maximum.reduce propagates NaN; images here are finite
      dval[(r * (W - 1) + c) * 2] = std::max(std::max(a, b), d);
      dval[(r * (W - 1) + c) * 2 + 1] = std::max(std::max(b, d), e);
    }
  dval[n_dual] = std::numeric_limits<float>::infinity();
  const float atol = (float)1e-8;
  std::vector<char> nonzero(n_verts);
  for (int i = 0; i < n_verts; ++i) nonzero[i] = !(std::fabs(val[i]) <= atol);

  std::vector<DmtEdge> E;
  E.reserve((size_t)(H - 1) * W + (size_t)H * (W - 1) + (size_t)(H - 1) * (W - 1));
  auto push = [&](int v1, int v2, int dv1, int dv2) {
    if (nonzero[v1] && nonzero[v2])
      E.push_back({v1, v2, dv1, dv2, 0, std::numeric_limits<float>::infinity(), std::max(val[v1], val[v2])});
  };
  for (int r = 0; r + 1 < H; ++r)                           // vertical
    for (int c = 0; c < W; ++c) {
      const int di = r * 2 * (W - 1) + c * 2;
      push(r * W + c, r * W + c + W, c == 0 ? n_dual : di - 1, c == W - 1 ? n_dual : di);
    }
  for (int r = 0; r < H; ++r)                               // horizontal
    for (int c = 0; c + 1 < W; ++c) {
      const int di = r * 2 * (W - 1) + c * 2;
      push(r * W + c, r * W + c + 1, r == 0 ? n_dual : di - 2 * (W - 1) + 1, r == H - 1 ? n_dual : di);
    }
  for (int r = 0; r + 1 < H; ++r)                           // diagonal
    for (int c = 0; c + 1 < W; ++c) {
      const int di = r * 2 * (W - 1) + c * 2;
      push(r * W + c + 1, r * W + c + 1 + W - 1, di, di + 1);
    }
  // np.lexsort((index, max_val)): ascending value, creation order among equals
  std::stable_sort(E.begin(), E.end(), [](const DmtEdge& a, const DmtEdge& b) { return a.max_val < b.max_val; });

  std::vector<int> parent(n_verts);
  std::iota(parent.begin(), parent.end(), 0);
  for (DmtEdge& e : E) {
    const float birth = uf_merge_host(parent, val, e.v1, e.v2, false);
    if (!std::isnan(birth)) { e.persistence = e.max_val - birth; e.pair_type = 1; }
  }
  std::reverse(E.begin(), E.end());
  std::vector<int> dparent(n_dual + 1);
  std::iota(dparent.begin(), dparent.end(), 0);
  for (DmtEdge& e : E) {
    if (e.pair_type != 0) continue;
    const float death = uf_merge_host(dparent, dval, e.dv1, e.dv2, true);
    if (!std::isnan(death)) { e.persistence = death - e.max_val; e.pair_type = 2; }
  }

  // ---- collect (:330-453) ----
  std::vector<int> nb((size_t)n_verts * 4, -1);
  for (const DmtEdge& e : E) {
    if (e.pair_type == 1 && e.persistence < delta1) {
      const int pr[2][2] = {{e.v1, e.v2}, {e.v2, e.v1}};
      for (auto& p : pr)
        for (int k = 0; k < 4; ++k)
          if (nb[(size_t)p[0] * 4 + k] == -1) { nb[(size_t)p[0] * 4 + k] = p[1]; break; }
    }
  }
  std::vector<int> morse_parent(n_verts, -1), explored(n_verts, -1), queue;
  queue.reserve(n_verts);
  for (int i = 0; i < n_verts; ++i) {
    if (morse_parent[i] != -1) continue;
    queue.clear();
    queue.push_back(i);
    int min_v = i;
    for (size_t qs = 0; qs < queue.size(); ++qs) {
      const int cur = queue[qs];
      explored[cur] = i;
      if (val[cur] < val[min_v] || (val[cur] == val[min_v] && cur < min_v)) min_v = cur;
      for (int k = 0; k < 4; ++k) {
        const int nbr = nb[(size_t)cur * 4 + k];
        if (nbr == -1) break;
        if (explored[nbr] != i) queue.push_back(nbr);
      }
    }
    morse_parent[min_v] = min_v;
    queue.clear();
    queue.push_back(min_v);
    for (size_t qs = 0; qs < queue.size(); ++qs) {
      const int cur = queue[qs];
      for (int k = 0; k < 4; ++k) {
        const int nbr = nb[(size_t)cur * 4 + k];
        if (nbr == -1) break;
        if (morse_parent[nbr] == -1) { morse_parent[nbr] = cur; queue.push_back(nbr); }
      }
    }
  }
  std::vector<char> in_mun(n_verts, 0);
  std::vector<std::pair<int, int>> mun;
  auto path_to_min = [&](int v) {
    int cur = v;
    while (!in_mun[cur] && morse_parent[cur] != cur && morse_parent[cur] != -1) {
      in_mun[cur] = 1;
      mun.emplace_back(cur, morse_parent[cur]);
      cur = morse_parent[cur];
    }
  };
  const float neg_d2 = -delta2;
  for (const DmtEdge& e : E) {
    if (e.persistence > delta1 && e.max_val < neg_d2) {
      path_to_min(e.v1);
      path_to_min(e.v2);
      mun.emplace_back(e.v1, e.v2);
    }
  }
  std::vector<int> index_map(n_verts, -1);
  std::vector<int32_t> vpos, eidx;
  eidx.reserve(mun.size() * 2);
  for (const auto& me : mun) {
    for (int v : {me.first, me.second}) {
      if (index_map[v] == -1) {
        index_map[v] = (int)(vpos.size() / 2);
        vpos.push_back(v / W);          // X = row
        vpos.push_back(v % W);          // Y = column
      }
    }
    eidx.push_back(index_map[me.first]);
    eidx.push_back(index_map[me.second]);
  }
  *n_vertices = (int64_t)(vpos.size() / 2);
  *n_edges_out = (int64_t)(eidx.size() / 2);
  if (!vertices_host && !edges_host) return TMAT_OK;        // size query
  TMAT_REQUIRE(vertices_host && edges_host, "tmat_dmt_graph: pass both output arrays or neither");
  TMAT_REQUIRE(*n_vertices <= cap_vertices && *n_edges_out <= cap_edges,
               "tmat_dmt_graph: capacity too small: need %lld vertices and %lld edges", (long long)*n_vertices,
               (long long)*n_edges_out);
  if (!vpos.empty()) memcpy(vertices_host, vpos.data(), vpos.size() * sizeof(int32_t));
  if (!eidx.empty()) memcpy(edges_host, eidx.data(), eidx.size() * sizeof(int32_t));
  return TMAT_OK;
}
