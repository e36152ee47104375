// bvh_build.hpp — host builder of the flat BVH2 handed to the GPU (K9 of SURVEY.md §2.1).
// Replaces the reference's random-axis median split pointer tree (bvh.hpp:15-44) by a binned-SAH
// build over double-precision boxes, emitted as 64-byte two-child nodes (zb_bvh_node) in DFS order
// with the child boxes rounded OUTWARD to float so the fp32 slab test stays conservative.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <numeric>
#include <vector>

#include "../../include/zenith_b200.h"
#include "zenith/math.hpp"

namespace zenith {

struct build_box {
    double lo[3], hi[3];
    void grow(const build_box& b) {
        for (int k = 0; k < 3; k++) { lo[k] = std::fmin(lo[k], b.lo[k]); hi[k] = std::fmax(hi[k], b.hi[k]); }
    }
    static build_box empty() { return {{infinity, infinity, infinity}, {-infinity, -infinity, -infinity}}; }
    double half_area() const {
        double d[3] = {hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2]};
        if (d[0] < 0 || d[1] < 0 || d[2] < 0) return 0;
        return d[0] * d[1] + d[1] * d[2] + d[2] * d[0];
    }
    static build_box from(const aabb& b) { return {{b.x.min, b.y.min, b.z.min}, {b.x.max, b.y.max, b.z.max}}; }
};

inline float round_down(double x) {
    float f = float(x);
    return double(f) > x ? std::nextafterf(f, -INFINITY) : f;
}
inline float round_up(double x) {
    float f = float(x);
    return double(f) < x ? std::nextafterf(f, INFINITY) : f;
}

class bvh_builder {
public:
    // boxes: one per primitive.  On return `order[i]` is the primitive stored at leaf slot i and
    // `nodes` holds the tree (root = nodes[0]); leaves reference [first, first+count) of `order`.
    // leaf_base is added to every leaf `first` (so several BLASes can share one primitive array).
    // max_depth: cap on the number of inner nodes along any root-to-leaf path (ZB_MAX_TLAS_DEPTH / ZB_MAX_BLAS_DEPTH;
    // the device traversal stack holds one deferred child per level).  SAH splits are taken while both children can
    // still be finished by balanced median splits within the cap; from there on the subtree is split by count.
    void build(const std::vector<build_box>& boxes, int max_leaf, uint32_t leaf_base, std::vector<zb_bvh_node>& nodes,
               std::vector<uint32_t>& order, int max_depth = ZB_MAX_BLAS_DEPTH) {
        in = &boxes;
        depth_cap = std::max(1, max_depth);
        leaf_cap = std::max(1, std::min(max_leaf, 8));
        base = leaf_base;
        const size_t n = boxes.size();
        order.resize(n);
        std::iota(order.begin(), order.end(), 0u);
        tmp.clear();
        nodes.clear();
        if (n == 0) return;
        cent.resize(n * 3);
        for (size_t i = 0; i < n; i++)
            for (int k = 0; k < 3; k++) cent[3 * i + k] = 0.5 * (boxes[i].lo[k] + boxes[i].hi[k]);
        ord = &order;
        int root = split(0, n, 0);
        if (tmp[root].count > 0) {  // single-leaf scene: wrap it so the root is always an inner node
            tnode r;
            r.box = tmp[root].box;
            r.left = root;
            r.right = -1;
            r.first = 0;
            r.count = 0;
            tmp.push_back(r);
            root = int(tmp.size()) - 1;
        }
        nodes.reserve(tmp.size());
        emit(root, nodes);
    }

private:
    struct tnode {
        build_box box;
        int left = -1, right = -1;
        size_t first = 0;
        uint32_t count = 0;  // >0: leaf
    };
    const std::vector<build_box>* in = nullptr;
    std::vector<uint32_t>* ord = nullptr;
    std::vector<double> cent;
    std::vector<tnode> tmp;
    int leaf_cap = 4;
    int depth_cap = ZB_MAX_BLAS_DEPTH;
    // inner levels a balanced (split-by-count) tree over n primitives needs, this node included; 0 = one leaf
    int balanced_levels(size_t n) const {
        int l = 0;
        while (n > size_t(leaf_cap)) { n = (n + 1) / 2; l++; }
        return l;
    }
public:
    // cost of one primitive test in units of one node visit (SAH leaf-versus-split decision)
    double isect_cost = 1.0;
private:
    uint32_t base = 0;

    int make_leaf(size_t lo, size_t hi, const build_box& box) {
        tnode t;
        t.box = box;
        t.first = lo;
        t.count = uint32_t(hi - lo);
        tmp.push_back(t);
        return int(tmp.size()) - 1;
    }

    int split(size_t lo, size_t hi, int depth) {  // depth = inner nodes above this one
        auto& o = *ord;
        const size_t n = hi - lo;
        build_box box = build_box::empty(), cbox = build_box::empty();
        for (size_t i = lo; i < hi; i++) {
            box.grow((*in)[o[i]]);
            for (int k = 0; k < 3; k++) {
                cbox.lo[k] = std::fmin(cbox.lo[k], cent[3 * o[i] + k]);
                cbox.hi[k] = std::fmax(cbox.hi[k], cent[3 * o[i] + k]);
            }
        }
        if (n == 1) return make_leaf(lo, hi, box);
        constexpr int NB = 16;
        const double c_trav = 1.0, c_isect = isect_cost;
        double best = infinity;
        int best_axis = -1, best_bin = -1;
        for (int axis = 0; axis < 3; axis++) {
            double ext = cbox.hi[axis] - cbox.lo[axis];
            if (!(ext > 0)) continue;
            build_box bb[NB];
            size_t cnt[NB] = {0};
            for (int b = 0; b < NB; b++) bb[b] = build_box::empty();
            double k1 = NB * (1 - 1e-9) / ext;
            for (size_t i = lo; i < hi; i++) {
                int b = int(k1 * (cent[3 * o[i] + axis] - cbox.lo[axis]));
                b = std::max(0, std::min(NB - 1, b));
                cnt[b]++;
                bb[b].grow((*in)[o[i]]);
            }
            double la[NB - 1], ra[NB - 1];
            size_t lc[NB - 1], rc[NB - 1];
            build_box acc = build_box::empty();
            size_t c = 0;
            for (int b = 0; b < NB - 1; b++) { acc.grow(bb[b]); c += cnt[b]; la[b] = acc.half_area(); lc[b] = c; }
            acc = build_box::empty();
            c = 0;
            for (int b = NB - 1; b > 0; b--) { acc.grow(bb[b]); c += cnt[b]; ra[b - 1] = acc.half_area(); rc[b - 1] = c; }
            for (int b = 0; b < NB - 1; b++) {
                if (lc[b] == 0 || rc[b] == 0) continue;
                double cost = la[b] * double(lc[b]) + ra[b] * double(rc[b]);
                if (cost < best) { best = cost; best_axis = axis; best_bin = b; }
            }
        }
        const double leaf_cost = c_isect * double(n);
        const double ha = box.half_area();
        const double split_cost = ha > 0 ? c_trav + c_isect * best / ha : infinity;
        if (n <= size_t(leaf_cap) && (best_axis < 0 || leaf_cost <= split_cost || depth >= depth_cap)) return make_leaf(lo, hi, box);
        size_t mid;
        bool by_count = best_axis < 0;  // all centroids coincide: split by count
        if (!by_count) {
            // would the SAH split leave a child that cannot be finished within the depth cap?
            size_t nl = 0;
            {
                double ext = cbox.hi[best_axis] - cbox.lo[best_axis];
                double k1 = NB * (1 - 1e-9) / ext;
                for (size_t i = lo; i < hi; i++) {
                    int b = int(k1 * (cent[3 * o[i] + best_axis] - cbox.lo[best_axis]));
                    nl += std::max(0, std::min(NB - 1, b)) <= best_bin ? 1 : 0;
                }
            }
            if (nl == 0 || nl == n) nl = n / 2;
            if (depth + 1 + std::max(balanced_levels(nl), balanced_levels(n - nl)) > depth_cap) by_count = true;
        }
        if (by_count) {
            mid = lo + (n + 1) / 2;
            int axis = 0;
            for (int k = 1; k < 3; k++)
                if (cbox.hi[k] - cbox.lo[k] > cbox.hi[axis] - cbox.lo[axis]) axis = k;
            if (best_axis >= 0)  // depth-capped: median along the widest centroid axis
                std::nth_element(o.begin() + lo, o.begin() + mid, o.begin() + hi,
                                 [&](uint32_t a, uint32_t b) { return cent[3 * a + axis] < cent[3 * b + axis]; });
        } else {
            double ext = cbox.hi[best_axis] - cbox.lo[best_axis];
            double k1 = NB * (1 - 1e-9) / ext;
            auto it = std::partition(o.begin() + lo, o.begin() + hi, [&](uint32_t p) {
                int b = int(k1 * (cent[3 * p + best_axis] - cbox.lo[best_axis]));
                b = std::max(0, std::min(NB - 1, b));
                return b <= best_bin;
            });
            mid = size_t(it - o.begin());
            if (mid == lo || mid == hi) mid = lo + n / 2;
        }
        int self = int(tmp.size());
        tmp.emplace_back();
        tmp[self].box = box;
        int l = split(lo, mid, depth + 1);
        int r = split(mid, hi, depth + 1);
        tmp[self].left = l;
        tmp[self].right = r;
        return self;
    }

    int32_t child_code(int t, std::vector<zb_bvh_node>& out) {
        if (t < 0) return ZB_CHILD_EMPTY;
        if (tmp[t].count > 0) {
            uint32_t code = (uint32_t(tmp[t].first + base) << 3) | (tmp[t].count - 1);
            return int32_t(~code);
        }
        return emit(t, out);
    }

    int32_t emit(int t, std::vector<zb_bvh_node>& out) {
        int32_t self = int32_t(out.size());
        out.emplace_back();
        const tnode tn = tmp[t];
        float l[6], r[6];
        pack(tn.left, l);
        pack(tn.right, r);
        zb_bvh_node nd;
        nd.lx_ly[0] = l[0]; nd.lx_ly[1] = l[3]; nd.lx_ly[2] = l[1]; nd.lx_ly[3] = l[4];
        nd.rx_ry[0] = r[0]; nd.rx_ry[1] = r[3]; nd.rx_ry[2] = r[1]; nd.rx_ry[3] = r[4];
        nd.lz_rz[0] = l[2]; nd.lz_rz[1] = l[5]; nd.lz_rz[2] = r[2]; nd.lz_rz[3] = r[5];
        nd.pad0 = nd.pad1 = 0;
        nd.left = child_code(tn.left, out);
        nd.right = child_code(tn.right, out);
        out[self] = nd;
        return self;
    }

    void pack(int t, float* b) const {  // lo.xyz, hi.xyz; inverted box for a missing child
        if (t < 0) {
            b[0] = b[1] = b[2] = INFINITY;
            b[3] = b[4] = b[5] = -INFINITY;
            return;
        }
        for (int k = 0; k < 3; k++) {
            b[k] = round_down(tmp[t].box.lo[k]);
            b[3 + k] = round_up(tmp[t].box.hi[k]);
        }
    }
};

}  // namespace zenith
