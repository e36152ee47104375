// wide_bvh.hpp — 8-wide compressed BVH (Ylitie, Karras, Laine 2017 layout) collapsed from the binned-SAH binary tree of
// bvh_build.hpp.  Replaces, for scenes with meshes, the traversal of the reference's pointer tree (bvh.hpp:46-54, :112-118):
// a third of the dependent node fetches per ray and 80 bytes per 8 children instead of 64 bytes per 2.
//
// Node = 80 bytes (zb_wide_node, include/zenith_b200.h): box origin p (3 floats), one power-of-two scale per axis (biased
// exponent bytes), 8 children with 8-bit quantised boxes lo = p + qlo 2^e, hi = p + qhi 2^e (rounded OUTWARDS, plus 1/32 of
// a step of guard for the device's fused decode), per-child meta byte (high 3 bits: number of primitives, 1 for an inner
// child, 0 = empty; low 5 bits: bit position — 24 + slot for inner children, primitive offset for leaves), imask (which
// slots hold inner children), child_base (inner children are consecutive nodes in slot order), prim_base (the leaf
// children's primitives are consecutive: at most 24 per node, at most 4 per leaf).
// Children sit in the slot whose sign vector (+-1, +-1, +-1) best matches their position relative to the node centre, so a
// ray visits slots in the order of (slot XOR octant) — front to back without sorting.
// Primitives are re-ordered so that every node's leaf primitives are consecutive; the leaf codes of the binary tree are
// remapped to the new order (both trees address the same primitive arrays).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/zenith_b200.h"

namespace zenith {

class wide_builder {
public:
    // nodes2: the binary tree (root = nodes2[0], child codes as in zb_bvh_node; leaf firsts are ABSOLUTE and lie in
    // [leaf_base, leaf_base + n_prims)).  On return: `wide` holds the 8-wide tree (root = wide[0], child_base LOCAL to
    // `wide`, prim_base absolute), perm[new_slot] = old_slot (both relative to leaf_base), the leaf codes of nodes2 point
    // at the new slots, depth = wide nodes on the longest root-to-leaf path.  Returns false when the binary tree has a
    // leaf of more than 4 primitives (the caller then ships no wide tree).
    bool build(std::vector<zb_bvh_node>& nodes2, uint32_t leaf_base, uint32_t n_prims, std::vector<zb_wide_node>& wide,
               std::vector<uint32_t>& perm, int& depth) {
        n2 = &nodes2;
        base = leaf_base;
        out = &wide;
        order = &perm;
        wide.clear();
        perm.clear();
        perm.reserve(n_prims);
        max_depth = 0;
        ok = true;
        if (nodes2.empty()) { depth = 0; return true; }
        wide.emplace_back();
        emit(0, 0, 1);
        depth = max_depth;
        return ok && perm.size() == n_prims;
    }

private:
    struct ent {
        float lo[3], hi[3];
        int32_t code;      // child code of the binary tree: >= 0 inner node, < 0 leaf
        int32_t from;      // binary node holding this child ...
        int side;          // ... as its left (0) or right (1) child
        bool closed;       // may not be opened further (constraints)
    };
    std::vector<zb_bvh_node>* n2 = nullptr;
    std::vector<zb_wide_node>* out = nullptr;
    std::vector<uint32_t>* order = nullptr;
    uint32_t base = 0;
    int max_depth = 0;
    bool ok = true;

    static double area(const ent& e) {
        double d[3] = {double(e.hi[0]) - e.lo[0], double(e.hi[1]) - e.lo[1], double(e.hi[2]) - e.lo[2]};
        if (d[0] < 0 || d[1] < 0 || d[2] < 0) return 0;
        return d[0] * d[1] + d[1] * d[2] + d[2] * d[0];
    }
    static uint32_t leaf_count(int32_t code) { return ((~uint32_t(code)) & 7u) + 1u; }
    static uint32_t leaf_first(int32_t code) { return (~uint32_t(code)) >> 3; }

    void children_of(int32_t node, std::vector<ent>& v) const {
        const zb_bvh_node& n = (*n2)[node];
        ent l{{n.lx_ly[0], n.lx_ly[2], n.lz_rz[0]}, {n.lx_ly[1], n.lx_ly[3], n.lz_rz[1]}, n.left, node, 0, false};
        ent r{{n.rx_ry[0], n.rx_ry[2], n.lz_rz[2]}, {n.rx_ry[1], n.rx_ry[3], n.lz_rz[3]}, n.right, node, 1, false};
        if (n.left != ZB_CHILD_EMPTY) v.push_back(l);
        if (n.right != ZB_CHILD_EMPTY) v.push_back(r);
    }

    void emit(int32_t node2, uint32_t w, int depth) {
        max_depth = std::max(max_depth, depth);
        std::vector<ent> e;
        e.reserve(9);
        children_of(node2, e);
        auto leaf_prims = [&]() {
            uint32_t t = 0;
            for (const ent& x : e)
                if (x.code < 0) t += leaf_count(x.code);
            return t;
        };
        // open the inner child of largest area while slots and the 24-primitive budget allow
        for (;;) {
            int best = -1;
            double best_area = -1;
            for (int i = 0; i < int(e.size()); i++)
                if (e[i].code >= 0 && !e[i].closed && area(e[i]) > best_area) { best = i; best_area = area(e[i]); }
            if (best < 0) break;
            std::vector<ent> c;
            children_of(e[best].code, c);
            uint32_t add = 0;
            for (const ent& x : c)
                if (x.code < 0) add += leaf_count(x.code);
            if (e.size() - 1 + c.size() > 8 || leaf_prims() + add > 24) { e[best].closed = true; continue; }
            e.erase(e.begin() + best);
            e.insert(e.end(), c.begin(), c.end());
        }
        // node box, origin and per-axis power-of-two scale
        double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (const ent& x : e)
            for (int k = 0; k < 3; k++) { lo[k] = std::fmin(lo[k], x.lo[k]); hi[k] = std::fmax(hi[k], x.hi[k]); }
        zb_wide_node nd;
        std::memset(&nd, 0, sizeof(nd));
        double scale[3];
        for (int k = 0; k < 3; k++) {
            float p = float(lo[k]);
            if (double(p) > lo[k]) p = std::nextafterf(p, -INFINITY);
            nd.p[k] = p;
            const double ext = std::fmax(hi[k] - double(p), 0.0);
            int eb = 1;  // biased exponent: scale = 2^(eb - 127) >= ext / 254
            if (ext > 0) eb = std::max(1, std::min(239, int(std::ceil(std::log2(ext / 254.0))) + 127));
            while (eb < 239 && std::ldexp(1.0, eb - 127) * 254.0 < ext) eb++;
            nd.e[k] = uint8_t(eb);
            scale[k] = std::ldexp(1.0, eb - 127);
        }
        // slots: child c goes where sum of dot(centre_c - centre_node, sign vector of the slot) is greedily largest
        int slot_of[8], child_in[8];
        std::fill(slot_of, slot_of + 8, -1);
        std::fill(child_in, child_in + 8, -1);
        {
            struct pair_cost { double cost; int c, s; };
            std::vector<pair_cost> pc;
            for (int c = 0; c < int(e.size()); c++)
                for (int s = 0; s < 8; s++) {
                    double cost = 0;
                    for (int k = 0; k < 3; k++) {
                        const double rel = 0.5 * (double(e[c].lo[k]) + e[c].hi[k]) - 0.5 * (lo[k] + hi[k]);
                        cost += ((s >> (2 - k)) & 1) ? rel : -rel;  // bit 2 = x, bit 1 = y, bit 0 = z; set = towards +
                    }
                    pc.push_back({cost, c, s});
                }
            std::stable_sort(pc.begin(), pc.end(), [](const pair_cost& a, const pair_cost& b) { return a.cost > b.cost; });
            for (const pair_cost& x : pc)
                if (slot_of[x.c] < 0 && child_in[x.s] < 0) { slot_of[x.c] = x.s; child_in[x.s] = x.c; }
        }
        // children in slot order: inner ones get consecutive node indices, leaf ones consecutive primitive slots
        uint32_t n_inner = 0;
        for (const ent& x : e) n_inner += x.code >= 0 ? 1u : 0u;
        const uint32_t child_base = uint32_t(out->size());
        out->resize(out->size() + n_inner);
        nd.child_base = child_base;
        nd.prim_base = base + uint32_t(order->size());
        uint32_t prim_off = 0, inner_seen = 0;
        struct todo { int32_t node2; uint32_t w; };
        std::vector<todo> next;
        for (int s = 0; s < 8; s++) {
            nd.qlo[0][s] = nd.qlo[1][s] = nd.qlo[2][s] = 255;  // empty slot: inverted box, no bits
            nd.qhi[0][s] = nd.qhi[1][s] = nd.qhi[2][s] = 0;
            const int c = child_in[s];
            if (c < 0) continue;
            const ent& x = e[c];
            for (int k = 0; k < 3; k++) {
                const double ql = std::floor((double(x.lo[k]) - nd.p[k]) / scale[k] - 1.0 / 32.0);
                const double qh = std::ceil((double(x.hi[k]) - nd.p[k]) / scale[k] + 1.0 / 32.0);
                nd.qlo[k][s] = uint8_t(std::max(0.0, std::min(255.0, ql)));
                nd.qhi[k][s] = uint8_t(std::max(0.0, std::min(255.0, qh)));
            }
            if (x.code >= 0) {
                nd.imask |= uint8_t(1u << s);
                nd.meta[s] = uint8_t((1u << 5) | (24u + uint32_t(s)));
                next.push_back({x.code, child_base + inner_seen});
                inner_seen++;
            } else {
                const uint32_t cnt = leaf_count(x.code), first = leaf_first(x.code);
                if (cnt > 4 || prim_off + cnt > 24) { ok = false; continue; }
                nd.meta[s] = uint8_t((cnt << 5) | prim_off);
                const uint32_t new_first = base + uint32_t(order->size());
                for (uint32_t j = 0; j < cnt; j++) order->push_back(first - base + j);
                prim_off += cnt;
                // the binary tree's leaf now lives at the new position
                int32_t& slot2 = x.side ? (*n2)[x.from].right : (*n2)[x.from].left;
                slot2 = int32_t(~((new_first << 3) | (cnt - 1)));
            }
        }
        (*out)[w] = nd;
        for (const todo& t : next) emit(t.node2, t.w, depth + 1);
    }
};

}  // namespace zenith
