/*
 * uavmvs_oracle.cpp -- CPU restatement of uavmvs's visibility-and-
 * reconstructability evaluation.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker (or the
 * timed CPU baseline).  The product (uavmvs_b200/csrc) never links, imports or
 * executes anything in oracle/.
 *
 * Parity status: the reference ships no tests, golden vectors or known-answer
 * fixtures for this path, so parity is pinned against
 * outputs of the reference's own code run here instead: the tree steps (SAH
 * build, ray traversal, KD nearest neighbour) are checked against
 * oracle/_ref/libref_acc.so, which is the unmodified libs/acc headers compiled
 * where they lie (tests/test_oracle_vs_ref.py, tests/golden/ made by
 * tests/golden/make_golden.py).  The kernel math of libs/eval/kernels.cu has no
 * CPU implementation in the reference (CUDA only, does not run in the build
 * container): that part is "parity unpinned" and is a line-by-line
 * restatement of the cited CUDA source with IEEE arithmetic.
 *
 * Arithmetic contract: fp32, round-to-nearest, no FMA
 * contraction (-ffp-contract=off), IEEE division and sqrt, no flush-to-zero.
 * Every function cites the reference file:line (relative to /root/reference)
 * it restates.
 */
#include <omp.h>
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <map>
#include <numeric>
#include <random>
#include <utility>
#include <vector>

namespace {

typedef uint32_t u32;
const u32 NAI = 0xffffffffu;
const float INF = std::numeric_limits<float>::infinity();

struct V3 { float x, y, z; };

inline float get(V3 const & a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
inline V3 sub(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 add(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 mul(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
inline V3 divs(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
inline V3 neg(V3 a) { return V3{-a.x, -a.y, -a.z}; }
/* libs/cacc/vector.h:189-198 and MVE's inner_product: zero accumulator first. */
inline float dot(V3 a, V3 b) {
    float r = 0.0f;
    r = r + a.x * b.x;
    r = r + a.y * b.y;
    r = r + a.z * b.z;
    return r;
}
inline float norm(V3 a) { return std::sqrt(dot(a, a)); }
/* libs/cacc/vector.h:217-224 */
inline V3 cross(V3 a, V3 b) {
    return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline V3 normalized(V3 a) { return divs(a, norm(a)); }

/* std::min / std::max NaN semantics (libs/acc/primitives.h:127-128):
 * std::min(a,b) = (b < a) ? b : a ; std::max(a,b) = (a < b) ? b : a. */
inline float smin(float a, float b) { return (b < a) ? b : a; }
inline float smax(float a, float b) { return (a < b) ? b : a; }

struct Box { float mn[3], mx[3]; };

inline Box empty_box() { Box b; for (int i = 0; i < 3; ++i) { b.mn[i] = INF; b.mx[i] = -INF; } return b; }
/* libs/acc/primitives.h:39-57 */
inline void grow(Box & a, Box const & b) {
    for (int i = 0; i < 3; ++i) { a.mn[i] = std::min(a.mn[i], b.mn[i]); a.mx[i] = std::max(a.mx[i], b.mx[i]); }
}
inline Box unite(Box const & a, Box const & b) { Box r = a; grow(r, b); return r; }
/* libs/acc/primitives.h:83-90 */
inline float surface_area(Box const & b) {
    float e0 = b.mx[0] - b.mn[0];
    float e1 = b.mx[1] - b.mn[1];
    float e2 = b.mx[2] - b.mn[2];
    return 2.0f * (e0 * e1 + e1 * e2 + e2 * e0);
}
/* libs/acc/primitives.h:108-112 */
inline float mid(Box const & b, int d) { return (b.mn[d] + b.mx[d]) / 2.0f; }

struct Node { u32 first, last, left, right; Box box; };

struct Tree {
    std::vector<Node> nodes;
    std::vector<float> tris;   /* 9 floats per triangle, leaf order */
    std::vector<u32> indices;  /* original face ids */
};

/* ------------------------------------------------------------------ */
/* SAH builder: libs/acc/bvh_tree.h:160-399                            */
/* ------------------------------------------------------------------ */
const int NUM_BINS = 64;  /* BVHTREE_NUM_BINS, libs/acc/bvh_tree.h:26-28 */

struct Builder {
    std::vector<Node> nodes;
    std::vector<u32> idx;
    std::vector<Box> const * tri_boxes;
    std::atomic<u32> count;

    u32 new_node(u32 first, u32 last) {       /* :72-82 */
        u32 id = count++;
        Node & n = nodes[id];
        n.first = first; n.last = last; n.left = NAI; n.right = NAI;
        n.box = empty_box();
        return id;
    }

    int bin_of(Box const & b, int d, float lo, float hi) const {
        /* :199, :240 -- float -> char conversion exactly as the reference writes it
         * (flat axis gives NaN -> 0 on x86-64). */
        char k = ((mid(b, d) - lo) / (hi - lo)) * (NUM_BINS - 1);
        return (int)k;
    }

    /* Binned SAH split, :178-274 */
    bool split_binned(u32 id) {
        std::vector<Box> const & boxes = *tri_boxes;
        Node & node = nodes[id];
        u32 n = node.last - node.first;
        Box bin_box[NUM_BINS]; u32 bin_n[NUM_BINS]; Box right[NUM_BINS];
        std::vector<u32> bin(n);

        float min_cost = INF;
        u32 best_d = 0; char best_k = 0;
        for (char d = 0; d < 3; ++d) {
            float lo = node.box.mn[(int)d], hi = node.box.mx[(int)d];
            for (int k = 0; k < NUM_BINS; ++k) { bin_n[k] = 0; bin_box[k] = empty_box(); }
            for (std::size_t i = node.first; i < node.last; ++i) {
                Box const & b = boxes[idx[i]];
                char k = ((mid(b, d) - lo) / (hi - lo)) * (NUM_BINS - 1);
                grow(bin_box[(int)k], b);
                bin_n[(int)k] += 1;
                bin[i - node.first] = k;
            }
            right[NUM_BINS - 1] = bin_box[NUM_BINS - 1];
            for (std::size_t k = NUM_BINS - 1; k > 0; --k) right[k - 1] = unite(bin_box[k - 1], right[k]);

            Box left = bin_box[0];
            std::size_t nl = bin_n[0];
            for (std::size_t k = 1; k < (std::size_t)NUM_BINS; ++k) {
                std::size_t nr = n - nl;
                float cost = 1.2f + (surface_area(left) / surface_area(node.box) * nl
                    + surface_area(right[k]) / surface_area(node.box) * nr);
                if (cost <= min_cost) {
                    min_cost = cost;
                    /* the reference stores (d, idx) into pair<IdxType, char> */
                    best_d = (u32)d; best_k = (char)k;
                }
                nl += bin_n[k];
                grow(left, bin_box[k]);
            }
        }
        if (min_cost >= n) return false;

        char d = (char)best_d;
        u32 sidx = (u32)best_k;
        float lo = node.box.mn[(int)d], hi = node.box.mx[(int)d];
        for (int k = 0; k < NUM_BINS; ++k) { bin_n[k] = 0; bin_box[k] = empty_box(); }
        for (std::size_t i = node.first; i < node.last; ++i) {
            Box const & b = boxes[idx[i]];
            char k = ((mid(b, d) - lo) / (hi - lo)) * (NUM_BINS - 1);
            grow(bin_box[(int)k], b);
            bin_n[(int)k] += 1;
            bin[i - node.first] = k;
        }
        u32 l = node.first, r = node.last - 1;
        while (l < r) {
            if (bin[l - node.first] < sidx) { l += 1; continue; }
            if (bin[r - node.first] >= sidx) { r -= 1; continue; }
            std::swap(bin[l - node.first], bin[r - node.first]);
            std::swap(idx[l], idx[r]);
        }
        std::size_t m = bin[(l & r) - node.first] >= sidx ? (l & r) : (l & r) + 1;
        u32 first = node.first, last = node.last;
        u32 lc = new_node(first, (u32)m);
        u32 rc = new_node((u32)m, last);
        nodes[id].left = lc; nodes[id].right = rc;
        for (u32 k = 0; k < (u32)NUM_BINS; ++k) {
            if (k < sidx) grow(nodes[lc].box, bin_box[k]);
            else grow(nodes[rc].box, bin_box[k]);
        }
        return true;
    }

    struct MidLess {
        std::vector<Box> const * boxes; int d;
        bool operator()(u32 a, u32 b) const {
            float ma = mid((*boxes)[a], d), mb = mid((*boxes)[b], d);
            return ma < mb || (ma == mb && a < b);
        }
    };

    /* Exact sweep SAH split, :277-333 */
    bool split_sweep(u32 id) {
        std::vector<Box> const & boxes = *tri_boxes;
        Node & node = nodes[id];
        u32 n = node.last - node.first;
        float min_cost = INF;
        int best_d = 0; u32 best_i = 0;
        std::vector<Box> right(n);
        for (char d = 0; d < 3; ++d) {
            MidLess less = {&boxes, (int)d};
            std::sort(idx.begin() + node.first, idx.begin() + node.last, less);
            right[n - 1] = boxes[idx[node.last - 1]];
            for (u32 i = node.last - 1; i > node.first; --i)
                right[i - 1 - node.first] = unite(boxes[idx[i - 1]], right[i - node.first]);
            node.box = right[0];

            Box left = boxes[idx[node.first]];
            for (u32 i = node.first + 1; i < node.last; ++i) {
                u32 nl = i - node.first;
                u32 nr = n - nl;
                float cost = 1.2f + (surface_area(left) / surface_area(node.box) * nl
                    + surface_area(right[nl]) / surface_area(node.box) * nr);
                if (cost <= min_cost) { min_cost = cost; best_d = d; best_i = i; }
                grow(left, boxes[idx[i]]);
            }
        }
        if (min_cost >= n) return false;
        MidLess less = {&boxes, best_d};
        std::sort(idx.begin() + node.first, idx.begin() + node.last, less);
        u32 first = node.first, last = node.last;
        u32 lc = new_node(first, best_i);
        u32 rc = new_node(best_i, last);
        nodes[id].left = lc; nodes[id].right = rc;
        return true;
    }

    bool split_node(u32 id) {                 /* :160-175 */
        u32 n = nodes[id].last - nodes[id].first;
        return (n > (u32)NUM_BINS) ? split_binned(id) : split_sweep(id);
    }

    void split_subtree(u32 root) {            /* :144-156 (sequential branch) */
        std::vector<u32> stack(1, root);
        while (!stack.empty()) {
            u32 id = stack.back(); stack.pop_back();
            if (split_node(id)) { stack.push_back(nodes[id].left); stack.push_back(nodes[id].right); }
        }
    }

    /* The reference fans out with std::thread (:132-143); the resulting topology
     * does not depend on the fan-out because every split only reads/writes its
     * own [first,last) range and node ids are re-assigned breadth first below. */
    void split_parallel(u32 root, int depth) {
        if (depth <= 0) { split_subtree(root); return; }
        if (!split_node(root)) return;
        u32 l = nodes[root].left, r = nodes[root].right;
        #pragma omp task shared(nodes)
        split_parallel(l, depth - 1);
        split_parallel(r, depth - 1);
        #pragma omp taskwait
    }
};

Tree * build_tree(float const * verts, u32 const * faces, u32 num_faces, int threads)
{
    /* :335-363 */
    std::vector<Box> boxes(num_faces);
    std::vector<float> ttris((std::size_t)num_faces * 9);
    Builder b;
    b.nodes.resize(2 * (std::size_t)num_faces - 1);
    b.count = 0;
    b.tri_boxes = &boxes;
    u32 root = b.new_node(0, num_faces);
    for (u32 i = 0; i < num_faces; ++i) {
        for (int k = 0; k < 3; ++k) {
            float const * p = verts + 3 * (std::size_t)faces[3 * (std::size_t)i + k];
            ttris[9 * (std::size_t)i + 3 * k + 0] = p[0];
            ttris[9 * (std::size_t)i + 3 * k + 1] = p[1];
            ttris[9 * (std::size_t)i + 3 * k + 2] = p[2];
        }
        float const * t = &ttris[9 * (std::size_t)i];
        for (int k = 0; k < 3; ++k) {   /* libs/acc/primitives.h:74-81 */
            boxes[i].mn[k] = std::min(t[k], std::min(t[3 + k], t[6 + k]));
            boxes[i].mx[k] = std::max(t[k], std::max(t[3 + k], t[6 + k]));
        }
        grow(b.nodes[root].box, boxes[i]);
    }
    b.idx.resize(num_faces);
    std::iota(b.idx.begin(), b.idx.end(), 0u);

    if (threads > 1) {
        int depth = 0; while ((1 << depth) < 4 * threads) ++depth;
        #pragma omp parallel num_threads(threads)
        #pragma omp single
        b.split_parallel(root, depth);
    } else {
        b.split_subtree(root);
    }
    u32 total = b.count;

    /* Breadth-first re-order, :365-393 */
    Tree * tree = new Tree;
    tree->nodes.resize(total);
    std::vector<u32> nidx; nidx.reserve(num_faces);
    u32 out = 0, cur = 0, last = 0;
    tree->nodes[out++] = b.nodes[root];
    while (cur < total) {
        Node & node = tree->nodes[cur++];
        if (node.left != NAI && node.right != NAI) {
            tree->nodes[out] = b.nodes[node.left]; node.left = out++;
            tree->nodes[out] = b.nodes[node.right]; node.right = out++;
        } else {
            nidx.insert(nidx.end(), b.idx.begin() + node.first, b.idx.begin() + node.last);
            node.first = last;
            last = (u32)nidx.size();
            node.last = last;
        }
    }
    /* :395-398 */
    tree->indices.swap(nidx);
    tree->tris.resize((std::size_t)num_faces * 9);
    for (std::size_t i = 0; i < tree->indices.size(); ++i)
        std::memcpy(&tree->tris[9 * i], &ttris[9 * (std::size_t)tree->indices[i]], 36);
    return tree;
}

/* ------------------------------------------------------------------ */
/* Ray primitives: libs/acc/primitives.h:119-132, :134-153, :214-248   */
/* ------------------------------------------------------------------ */
struct Ray { V3 o, d; float tmin, tmax; };

inline bool hit_box(Ray const & ray, Box const & b, float * tmin_out)
{
    float tmin = ray.tmin, tmax = ray.tmax;
    for (int i = 0; i < 3; ++i) {
        float t1 = (b.mn[i] - get(ray.o, i)) / get(ray.d, i);
        float t2 = (b.mx[i] - get(ray.o, i)) / get(ray.d, i);
        tmin = smax(tmin, smin(smin(t1, t2), INF));
        tmax = smin(tmax, smax(smax(t1, t2), -INF));
    }
    *tmin_out = tmin;
    return tmax >= smax(tmin, 0.0f);
}

inline bool hit_tri(Ray const & ray, float const * t9, float * t_out)
{
    V3 a = {t9[0], t9[1], t9[2]}, b = {t9[3], t9[4], t9[5]}, c = {t9[6], t9[7], t9[8]};
    V3 ab = sub(b, a), ac = sub(c, a);
    V3 nrm = cross(ab, ac);
    float len = norm(nrm);
    if (len < FLT_EPSILON) return false;
    nrm = divs(nrm, len);
    float cosine = dot(nrm, ray.d);
    if (std::abs(cosine) < FLT_EPSILON) return false;
    float t = -dot(nrm, sub(ray.o, a)) / cosine;
    if (t < ray.tmin || ray.tmax < t) return false;
    V3 p = add(ray.o, mul(ray.d, t));
    V3 ap = sub(p, a);
    /* barycentric_coordinates(ab, ac, ap), :134-153 */
    float d00 = dot(ab, ab), d01 = dot(ab, ac), d11 = dot(ac, ac);
    float d20 = dot(ap, ab), d21 = dot(ap, ac);
    double denom = d00 * d11 - d01 * d01;
    float b1 = (d11 * d20 - d01 * d21) / denom;
    float b2 = (d00 * d21 - d01 * d20) / denom;
    float b0 = 1.0f - b1 - b2;
    const float eps = 1e-3f;
    if (-eps > b0 || b0 > 1.0f + eps) return false;
    if (-eps > b1 || b1 > 1.0f + eps) return false;
    if (-eps > b2 || b2 > 1.0f + eps) return false;
    *t_out = t;
    return true;
}


/* ------------------------------------------------------------------ */
/* closest point: libs/acc/primitives.h:155-163 (box), :174-212       */
/* (triangle), libs/acc/bvh_tree.h:469-548 (tree search)               */
/* ------------------------------------------------------------------ */
inline V3 closest_on_box(V3 v, Box const & b)
{
    float r[3];
    for (int i = 0; i < 3; ++i) r[i] = smax(b.mn[i], smin(get(v, i), b.mx[i]));
    return V3{r[0], r[1], r[2]};
}

inline V3 closest_on_tri(V3 vertex, float const * t9)
{
    V3 a = {t9[0], t9[1], t9[2]}, b = {t9[3], t9[4], t9[5]}, c = {t9[6], t9[7], t9[8]};
    V3 ab = sub(b, a), ac = sub(c, a);
    V3 normal = cross(ab, ac);
    float n = norm(normal);
    normal = divs(normal, n);
    V3 p = sub(vertex, mul(normal, dot(normal, sub(vertex, a))));
    V3 ap = sub(p, a);
    float d00 = dot(ab, ab), d01 = dot(ab, ac), d11 = dot(ac, ac);
    float d20 = dot(ap, ab), d21 = dot(ap, ac);
    double denom = d00 * d11 - d01 * d01;
    float b1 = (d11 * d20 - d01 * d21) / denom;
    float b2 = (d00 * d21 - d01 * d20) / denom;
    float b0 = 1.0f - b1 - b2;
    if (b0 < 0.0f) {
        V3 bc = sub(c, b);
        float len = norm(bc);
        float t = smax(0.0f, smin(dot(bc, sub(p, b)) / len, len));
        return add(b, mul(bc, t / len));
    }
    if (b1 < 0.0f) {
        V3 ca = sub(a, c);
        float len = norm(ca);
        float t = smax(0.0f, smin(dot(ca, sub(p, c)) / len, len));
        return add(c, mul(ca, t / len));
    }
    if (b2 < 0.0f) {
        float len = norm(ab);
        float t = smax(0.0f, smin(dot(ab, sub(p, a)) / len, len));
        return add(a, mul(ab, t / len));
    }
    return add(add(mul(a, b0), mul(b, b1)), mul(c, b2));
}

inline float sqdist(V3 a, V3 b) { V3 d = sub(a, b); return dot(d, d); }

inline bool tree_closest_point(Tree const & t, V3 vertex, float max_dist, u32 * idx_out, V3 * cp_out)
{
    float dist = max_dist * max_dist;
    u32 idx = NAI; V3 cp = {0.0f, 0.0f, 0.0f};
    u32 node_id = 0;
    std::vector<u32> s;
    while (true) {
        Node const & node = t.nodes[node_id];
        if (node.left != NAI && node.right != NAI) {
            float dl = sqdist(closest_on_box(vertex, t.nodes[node.left].box), vertex);
            float dr = sqdist(closest_on_box(vertex, t.nodes[node.right].box), vertex);
            bool left = dl < dist, right = dr < dist;
            if (left && right) {
                if (dl < dr) { s.push_back(node.right); node_id = node.left; }
                else { s.push_back(node.left); node_id = node.right; }
            } else {
                if (right) node_id = node.right;
                if (left) node_id = node.left;
            }
            if (!left && !right) {
                if (s.empty()) break;
                node_id = s.back(); s.pop_back();
            }
        } else {
            u32 idx_leaf = NAI; V3 cp_leaf = {0.0f, 0.0f, 0.0f}; float d_leaf = INF;
            for (u32 i = node.first; i < node.last; ++i) {
                V3 c = closest_on_tri(vertex, &t.tris[9 * (size_t)i]);
                float d = sqdist(c, vertex);
                if (d < d_leaf) { cp_leaf = c; d_leaf = d; idx_leaf = i; }
            }
            /* :530-536: the leaf's best is re-measured (NAI leaf result: cp is whatever, dist compares false for NaN/inf) */
            float dist_leaf = idx_leaf == NAI ? INF : sqdist(cp_leaf, vertex);
            if (dist_leaf < dist) { dist = dist_leaf; idx = idx_leaf; cp = cp_leaf; }
            if (s.empty()) break;
            node_id = s.back(); s.pop_back();
        }
    }
    if (idx == NAI) return false;
    *idx_out = idx; *cp_out = cp;
    return true;
}

struct Counters { uint64_t rays, inner, leaf, tri, hit; };

/* Traversal: libs/acc/bvh_tree.h:401-467.  any_hit stops at the first accepted
 * triangle (the boolean cannot change afterwards).  Otherwise closest hit with
 * ray.tmax shrinking (:450-452) and face id = indices[i] (:409). */
bool traverse(Tree const & tree, Ray ray, bool any_hit, u32 * face_out, float * t_out, Counters * c)
{
    float best_t = INF; u32 best_face = NAI;
    u32 stack[256]; int sp = 0;
    u32 id = 0;
    if (c) c->rays++;
    while (true) {
        Node const & node = tree.nodes[id];
        if (node.left != NAI && node.right != NAI) {
            if (c) c->inner++;
            float tl, tr;
            bool l = hit_box(ray, tree.nodes[node.left].box, &tl);
            bool r = hit_box(ray, tree.nodes[node.right].box, &tr);
            if (l && r) {
                if (tl < tr) { stack[sp++] = node.right; id = node.left; }
                else { stack[sp++] = node.left; id = node.right; }
            } else {
                if (r) id = node.right;
                if (l) id = node.left;
            }
            if (!l && !r) { if (sp == 0) break; id = stack[--sp]; }
        } else {
            if (c) c->leaf++;
            bool any = false;
            for (u32 i = node.first; i < node.last; ++i) {
                float t;
                if (c) c->tri++;
                if (hit_tri(ray, &tree.tris[9 * (std::size_t)i], &t)) {
                    if (t > best_t) continue;
                    if (c) c->hit++;
                    best_t = t; best_face = tree.indices[i]; any = true;
                    if (any_hit) { if (face_out) *face_out = best_face; if (t_out) *t_out = best_t; return true; }
                }
            }
            if (any) ray.tmax = best_t;
            if (sp == 0) break;
            id = stack[--sp];
        }
    }
    if (face_out) *face_out = best_face;
    if (t_out) *t_out = best_t;
    return best_t < INF;
}

/* ------------------------------------------------------------------ */
/* KD-tree over the sphere vertices                                    */
/* ------------------------------------------------------------------ */
struct KDNode { u32 vid, dim, left, right; };
struct KDTree { std::vector<KDNode> nodes; std::vector<V3> verts; };

/* libs/acc/kd_tree.h:92-158 with max_threads = 1 (breadth-first queue, :121-129):
 * node ids are then assigned in BFS order.  Median by std::nth_element on the
 * split axis, axis cycling x,y,z. */
KDTree * build_kd(float const * verts, u32 n)
{
    KDTree * t = new KDTree;
    t->verts.resize(n);
    for (u32 i = 0; i < n; ++i) t->verts[i] = V3{verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]};
    std::vector<u32> idx(n); std::iota(idx.begin(), idx.end(), 0u);
    struct Work { u32 d, first, last; };
    std::vector<Work> work; work.reserve(n);
    t->nodes.reserve(n);
    work.push_back(Work{0, 0, n});
    t->nodes.push_back(KDNode{NAI, 0, NAI, NAI});
    for (std::size_t q = 0; q < work.size(); ++q) {
        Work w = work[q];
        u32 m = (w.last + w.first) / 2;
        u32 d = w.d;
        std::vector<V3> const & vv = t->verts;
        std::nth_element(idx.begin() + w.first, idx.begin() + m, idx.begin() + w.last,
            [&vv, d](u32 a, u32 b) -> bool { return get(vv[a], d) < get(vv[b], d); });
        u32 nd = (d + 1) % 3;
        t->nodes[q].vid = idx[m];
        t->nodes[q].dim = d;
        if (m - w.first > 0) {
            t->nodes[q].left = (u32)t->nodes.size();
            t->nodes.push_back(KDNode{NAI, nd, NAI, NAI}); work.push_back(Work{nd, w.first, m});
        }
        if (w.last - (m + 1) > 0) {
            t->nodes[q].right = (u32)t->nodes.size();
            t->nodes.push_back(KDNode{NAI, nd, NAI, NAI}); work.push_back(Work{nd, m + 1, w.last});
        }
    }
    return t;
}

/* libs/cacc/nnsearch.h:67-155 (device search the kernels call): strict '<' so
 * the first vertex found at the minimum distance wins. */
bool kd_find_nn(KDTree const & kd, V3 q, u32 * idx_out, float * dist_out)
{
    u32 idx = NAI; float max_dist = INF;
    u32 node_idx = 0; bool down = true;
    u32 stack[128]; int sp = 0;
    while (true) {
        KDNode const & node = kd.nodes[node_idx];
        V3 vert = kd.verts[node.vid];
        float diff = get(q, node.dim) - get(vert, node.dim);
        if (down) {
            float dist = norm(sub(q, vert));
            if (dist < max_dist) { idx = node.vid; max_dist = dist; }
            if (node.left != NAI || node.right != NAI) {
                down = true;
                if (node.left != NAI && node.right != NAI) stack[sp++] = node_idx;
                u32 next = (diff < 0.0f) ? node.left : node.right;
                u32 other = (diff < 0.0f) ? node.right : node.left;
                node_idx = (next != NAI) ? next : other;
            } else { down = false; node_idx = NAI; }
        } else {
            if (std::abs(diff) < max_dist) { down = true; node_idx = (diff < 0.0f) ? node.right : node.left; }
            else { down = false; node_idx = NAI; }
        }
        if (node_idx == NAI) { if (sp == 0) break; node_idx = stack[--sp]; }
    }
    if (idx == NAI) return false;
    if (idx_out) *idx_out = idx;
    if (dist_out) *dist_out = max_dist;
    return true;
}

/* ------------------------------------------------------------------ */
/* Kernel math: libs/eval/kernels.cu                                   */
/* ------------------------------------------------------------------ */
const float PI = 3.14159265f;   /* :78 */

struct Heur { float m_k, m_x0, t_k, t_x0; };

/* :80-84 (device __expf restated with libm expf; tolerance 1e-5 rel) */
inline float logistic(float x, float k, float x0) { return 1.0f / (1.0f + std::exp(-k * (x - x0))); }
/* :87-91 (__powf(|x|, 2) restated as a square) */
inline float func(float recon, float target) {
    float a = std::abs(std::min(recon - target, 0.0f));
    return a * a;
}
/* :29-40 */
inline V3 orthogonal(V3 v) {
    V3 n0 = {1.0f, 0.0f, 0.0f}, n1 = {0.0f, 1.0f, 0.0f};
    if (std::abs(dot(n0, v)) < std::abs(dot(n1, v))) return cross(n0, v);
    return cross(n1, v);
}
/* :63-76 */
inline V3 relative_direction(V3 v2cn, V3 n) {
    V3 rx = normalized(orthogonal(n));
    V3 ry = normalized(cross(n, rx));
    V3 rz = n;
    V3 rel = {dot(rx, v2cn), dot(ry, v2cn), dot(rz, v2cn)};
    return normalized(rel);
}
/* :21-27 with Mat3f * Vec3f of libs/cacc/matrix.h:69-82 */
inline void project(V3 v, float const * K, float * px, float * py) {
    float p[3];
    for (int i = 0; i < 3; ++i) {
        float s = 0.0f;
        s += K[3 * i + 0] * v.x; s += K[3 * i + 1] * v.y; s += K[3 * i + 2] * v.z;
        p[i] = s;
    }
    *px = p[0] / p[2] - 0.5f;
    *py = p[1] / p[2] - 0.5f;
}
/* mult(w2c, v, 1.0f): libs/cacc/matrix.h:112-126 */
inline V3 mult_w2c(float const * W, V3 v) {
    float r[3];
    for (int i = 0; i < 3; ++i) {
        float s = 0.0f;
        s += W[4 * i + 0] * v.x; s += W[4 * i + 1] * v.y; s += W[4 * i + 2] * v.z;
        r[i] = s + W[4 * i + 3] * 1.0f;
    }
    return V3{r[0], r[1], r[2]};
}
/* :42-61, UNSAFE undefined */
inline bool visible(Tree const & tree, V3 v, V3 v2cn, float l, Counters * c) {
    Ray ray; ray.o = v; ray.d = v2cn; ray.tmin = l * 0.001f; ray.tmax = l;
    return !traverse(tree, ray, true, nullptr, nullptr, c);
}
/* :111-136; rows are float4 {rel.x, rel.y, rel.z, scale} at stride `stride` float4s */
inline float heuristic(float const * rows, std::size_t stride, u32 n, float const * nr, Heur const & h) {
    float sum = 0.0f;
    for (u32 i = 0; i < n; ++i) {
        float const * r = rows + 4 * stride * i;
        float calpha = dot(V3{nr[0], nr[1], nr[2]}, V3{r[0], r[1], r[2]});
        float alpha = std::acos(std::max(-1.0f, std::min(calpha, 1.0f)));
        float scale = std::min(r[3], nr[3]);
        float ctheta = std::min(r[2], nr[2]);
        float matchability = (1.0f - logistic(alpha, h.m_k, PI / h.m_x0)) * ctheta;
        float triangulation = logistic(alpha, h.t_k, PI / h.t_x0) * scale;
        sum += matchability * triangulation;
    }
    return sum;
}

inline V3 ld3(float const * p, std::size_t i) { return V3{p[3 * i], p[3 * i + 1], p[3 * i + 2]}; }

}  // namespace

extern "C" {

/* OpenMP team size of every parallel loop below.  torch.distributed.run exports OMP_NUM_THREADS=1 to its workers, which
 * would silently turn the CPU baseline into a single-thread run: bench.py sets the team explicitly and reports what it got. */
int orc_set_threads(int n)
{
    if (n > 0) omp_set_num_threads(n);
    int got = 1;
    #pragma omp parallel
    {
        #pragma omp single
        got = omp_get_num_threads();
    }
    return got;
}

/* ---- trees ---- */
void * orc_bvh_build(float const * verts, u32 num_verts, u32 const * faces, u32 num_faces, int threads)
{
    (void)num_verts;
    return build_tree(verts, faces, num_faces, threads);
}

void * orc_bvh_from_arrays(u32 const * nodes, float const * aabbs, float const * tris,
    u32 const * indices, u32 num_nodes, u32 num_tris)
{
    Tree * t = new Tree;
    t->nodes.resize(num_nodes);
    for (u32 i = 0; i < num_nodes; ++i) {
        Node & n = t->nodes[i];
        n.first = nodes[4 * i]; n.last = nodes[4 * i + 1]; n.left = nodes[4 * i + 2]; n.right = nodes[4 * i + 3];
        for (int k = 0; k < 3; ++k) { n.box.mn[k] = aabbs[6 * i + k]; n.box.mx[k] = aabbs[6 * i + 3 + k]; }
    }
    t->tris.assign(tris, tris + 9 * (std::size_t)num_tris);
    t->indices.assign(indices, indices + num_tris);
    return t;
}

u32 orc_bvh_num_nodes(void * h) { return (u32)((Tree *)h)->nodes.size(); }
u32 orc_bvh_num_tris(void * h) { return (u32)((Tree *)h)->indices.size(); }

void orc_bvh_export(void * h, u32 * nodes, float * aabbs, float * tris, u32 * indices)
{
    Tree const & t = *(Tree *)h;
    for (std::size_t i = 0; i < t.nodes.size(); ++i) {
        Node const & n = t.nodes[i];
        nodes[4 * i] = n.first; nodes[4 * i + 1] = n.last; nodes[4 * i + 2] = n.left; nodes[4 * i + 3] = n.right;
        for (int k = 0; k < 3; ++k) { aabbs[6 * i + k] = n.box.mn[k]; aabbs[6 * i + 3 + k] = n.box.mx[k]; }
    }
    std::memcpy(tris, t.tris.data(), t.tris.size() * 4);
    std::memcpy(indices, t.indices.data(), t.indices.size() * 4);
}

void orc_bvh_destroy(void * h) { delete (Tree *)h; }

/* rays: 8 floats each (origin, dir, tmin, tmax).  counters: 5 x u64 {rays, inner, leaf, tri, hit}. */
void orc_bvh_intersect(void * h, float const * rays, uint64_t n, int any_hit,
    uint8_t * hit, u32 * face, float * t, uint64_t * counters)
{
    Tree const & tree = *(Tree *)h;
    Counters total = {0, 0, 0, 0, 0};
    #pragma omp parallel
    {
        Counters c = {0, 0, 0, 0, 0};
        #pragma omp for schedule(dynamic, 1024)
        for (int64_t i = 0; i < (int64_t)n; ++i) {
            Ray ray;
            ray.o = ld3(rays + 8 * i, 0); ray.d = ld3(rays + 8 * i + 3, 0);
            ray.tmin = rays[8 * i + 6]; ray.tmax = rays[8 * i + 7];
            u32 f = NAI; float tt = 0.0f;
            bool r = traverse(tree, ray, any_hit != 0, &f, &tt, counters ? &c : nullptr);
            if (hit) hit[i] = r ? 1 : 0;
            if (face) face[i] = r ? f : NAI;
            if (t) t[i] = r ? tt : 0.0f;
        }
        #pragma omp critical
        { total.rays += c.rays; total.inner += c.inner; total.leaf += c.leaf; total.tri += c.tri; total.hit += c.hit; }
    }
    if (counters) { counters[0] = total.rays; counters[1] = total.inner; counters[2] = total.leaf; counters[3] = total.tri; counters[4] = total.hit; }
}

/* found[i] = 1/0, idx[i] = leaf-order triangle index (acc's cp.first) or NAI, cp = closest point */
void orc_bvh_closest_point(void * h, float const * points, uint64_t n, float max_dist,
    uint8_t * found, u32 * idx, float * cp)
{
    Tree const & t = *(Tree *)h;
    #pragma omp parallel for schedule(dynamic, 256)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        u32 id = NAI; V3 c = {0.0f, 0.0f, 0.0f};
        bool f = tree_closest_point(t, V3{points[3 * i], points[3 * i + 1], points[3 * i + 2]}, max_dist, &id, &c);
        found[i] = f ? 1 : 0; idx[i] = f ? id : NAI;
        cp[3 * i] = f ? c.x : 0.0f; cp[3 * i + 1] = f ? c.y : 0.0f; cp[3 * i + 2] = f ? c.z : 0.0f;
    }
}

void * orc_kd_build(float const * verts, u32 n) { return build_kd(verts, n); }
u32 orc_kd_num_nodes(void * h) { return (u32)((KDTree *)h)->nodes.size(); }
void * orc_kd_from_arrays(u32 const * nodes, float const * verts, u32 num_nodes, u32 num_verts)
{
    KDTree * t = new KDTree;
    t->nodes.resize(num_nodes);
    for (u32 i = 0; i < num_nodes; ++i) t->nodes[i] = KDNode{nodes[4 * i], nodes[4 * i + 1], nodes[4 * i + 2], nodes[4 * i + 3]};
    t->verts.resize(num_verts);
    for (u32 i = 0; i < num_verts; ++i) t->verts[i] = ld3(verts, i);
    return t;
}
void orc_kd_export(void * h, u32 * nodes, float * verts)
{
    KDTree const & t = *(KDTree *)h;
    for (std::size_t i = 0; i < t.nodes.size(); ++i) {
        nodes[4 * i] = t.nodes[i].vid; nodes[4 * i + 1] = t.nodes[i].dim; nodes[4 * i + 2] = t.nodes[i].left; nodes[4 * i + 3] = t.nodes[i].right;
    }
    for (std::size_t i = 0; i < t.verts.size(); ++i) { verts[3 * i] = t.verts[i].x; verts[3 * i + 1] = t.verts[i].y; verts[3 * i + 2] = t.verts[i].z; }
}
void orc_kd_find_nn(void * h, float const * q, uint64_t n, u32 * idx, float * dist)
{
    KDTree const & t = *(KDTree *)h;
    for (uint64_t i = 0; i < n; ++i) {
        u32 id = NAI; float d = 0.0f;
        kd_find_nn(t, ld3(q, i), &id, &d);
        idx[i] = id; if (dist) dist[i] = d;
    }
}
void orc_kd_destroy(void * h) { delete (KDTree *)h; }

/* libs/geom/sphere.h:17-115 -- icosahedron subdivided `subdivisions` times with
 * the reference's book-keeping (three midpoints pushed per face, first edge id
 * wins), so 3 subdivisions give 1272 vertices (642 referenced + 630 orphans).
 * Returns the vertex count; writes radius-scaled positions when out != NULL. */
u32 orc_sphere_vertices(float radius, u32 subdivisions, float * out, u32 capacity)
{
    static const float ico[12][3] = {
        {0.0f, -0.5257311f, 0.8506508f}, {0.0f, 0.5257311f, 0.8506508f},
        {0.0f, -0.5257311f, -0.8506508f}, {0.0f, 0.5257311f, -0.8506508f},
        {0.8506508f, 0.0f, 0.5257311f}, {0.8506508f, 0.0f, -0.5257311f},
        {-0.8506508f, 0.0f, 0.5257311f}, {-0.8506508f, 0.0f, -0.5257311f},
        {0.5257311f, 0.8506508f, 0.0f}, {0.5257311f, -0.8506508f, 0.0f},
        {-0.5257311f, 0.8506508f, 0.0f}, {-0.5257311f, -0.8506508f, 0.0f}};
    static const u32 ico_faces[60] = {
        0, 4, 1, 0, 9, 4, 9, 5, 4, 4, 5, 8, 4, 8, 1, 8, 10, 1, 8, 3, 10, 5, 3, 8, 5, 2, 3, 2, 7, 3,
        7, 10, 3, 7, 6, 10, 7, 11, 6, 11, 0, 6, 0, 1, 6, 6, 1, 10, 9, 0, 11, 9, 11, 2, 9, 2, 5, 7, 2, 11};
    std::vector<V3> verts;
    for (int i = 0; i < 12; ++i) verts.push_back(V3{ico[i][0], ico[i][1], ico[i][2]});
    std::vector<u32> faces(ico_faces, ico_faces + 60);
    for (u32 s = 0; s < subdivisions; ++s) {
        typedef std::pair<u32, u32> Edge;
        std::map<Edge, u32> edge_map;
        for (std::size_t j = 0; j < faces.size(); j += 3) {
            u32 v[3] = {faces[j], faces[j + 1], faces[j + 2]};
            for (int e = 0; e < 3; ++e) {
                u32 p = v[e], q = v[(e + 1) % 3];
                u32 id = (u32)verts.size();
                /* :70-78: both orientations map to the new id; insert keeps the first */
                edge_map.insert(std::make_pair(Edge(p, q), id));
                edge_map.insert(std::make_pair(Edge(q, p), id));
                verts.push_back(divs(add(verts[p], verts[q]), 2.0f));
            }
        }
        std::size_t nf = faces.size();
        for (std::size_t j = 0; j < nf; j += 3) {
            u32 v0 = faces[j], v1 = faces[j + 1], v2 = faces[j + 2];
            u32 e0 = edge_map[Edge(v0, v1)], e1 = edge_map[Edge(v1, v2)], e2 = edge_map[Edge(v2, v0)];
            u32 add9[9] = {e0, v1, e1, e1, v2, e2, e2, v0, e0};
            faces.insert(faces.end(), add9, add9 + 9);
            faces[j] = e0; faces[j + 1] = e1; faces[j + 2] = e2;
        }
    }
    if (out) {
        for (std::size_t i = 0; i < verts.size() && i < capacity; ++i) {
            V3 n = normalized(verts[i]);   /* :100-103 */
            V3 p = mul(n, radius);
            out[3 * i] = p.x; out[3 * i + 1] = p.y; out[3 * i + 2] = p.z;
        }
    }
    return (u32)verts.size();
}

/* ---- K1: update_observation_rays, libs/eval/kernels.cu:138-202 ----
 * One call = the reference's serial launch loop over `num_views` views
 * (apps/evaluate_trajectory/evaluate_trajectory.cu:154-164).  Points are
 * independent, so the loop nest is points (parallel) x views (in order); the
 * per-point row order equals the stream order of the reference.
 *   views: 28 floats per view {pos[3], w2c[16] row-major, calib[9] row-major}
 *   obs:   float4 rows, element (row r, point i) at obs[4*(r*stride + i)]
 *   state (optional, u8 [num_views * N]): 0 culled, 1 traced+occluded, 2 visible
 *   counters (optional): 5 x u64 traversal counters + [5] = pairs tested      */
void orc_update_observation_rays(void * bvh, int populate, float const * views, u32 num_views,
    float max_distance, int width, int height,
    float const * verts, float const * normals, u32 N,
    float * obs, u32 * num_rows, u32 max_rows, uint64_t stride,
    uint8_t * state, uint64_t * counters)
{
    Tree const & tree = *(Tree *)bvh;
    Counters total = {0, 0, 0, 0, 0};
    #pragma omp parallel
    {
        Counters c = {0, 0, 0, 0, 0};
        #pragma omp for schedule(dynamic, 256)
        for (int64_t id = 0; id < (int64_t)N; ++id) {
            V3 v = ld3(verts, id), n = ld3(normals, id);
            for (u32 k = 0; k < num_views; ++k) {
                float const * view = views + 28 * (std::size_t)k;
                if (state) state[(std::size_t)k * N + id] = 0;
                V3 view_pos = {view[0], view[1], view[2]};
                V3 v2c = sub(view_pos, v);
                float l = norm(v2c);
                V3 v2cn = divs(v2c, l);
                float ctheta = dot(v2cn, n);
                if (ctheta < 0.087f) continue;
                if (l >= max_distance) continue;
                float px, py;
                project(mult_w2c(view + 3, v), view + 19, &px, &py);
                if (px < 0.0f || width <= px || py < 0.0f || height <= py) continue;
                bool vis = visible(tree, v, v2cn, l, counters ? &c : nullptr);
                if (state) state[(std::size_t)k * N + id] = vis ? 2 : 1;
                if (!vis) continue;
                V3 rel = relative_direction(v2cn, n);
                float rel4[4] = {rel.x, rel.y, rel.z, 1.0f - (l / max_distance)};
                if (populate) {
                    u32 row = num_rows[id]; num_rows[id] = row + 1;
                    if (row >= max_rows) continue;
                    std::memcpy(obs + 4 * (row * stride + id), rel4, 16);
                } else {
                    u32 rows = std::min(num_rows[id], max_rows);
                    for (u32 i = 0; i < rows; ++i) {
                        float * o = obs + 4 * (i * stride + id);
                        bool equal = true;
                        for (int j = 0; j < 4; ++j) equal = equal && std::abs(rel4[j] - o[j]) < 1e-5f;
                        if (equal) { o[3] = -1.0f; break; }
                    }
                }
            }
        }
        #pragma omp critical
        { total.rays += c.rays; total.inner += c.inner; total.leaf += c.leaf; total.tri += c.tri; total.hit += c.hit; }
    }
    if (counters) {
        counters[0] = total.rays; counters[1] = total.inner; counters[2] = total.leaf;
        counters[3] = total.tri; counters[4] = total.hit; counters[5] = (uint64_t)N * num_views;
    }
}

/* ---- K2: process_observation_rays, :204-264 ----
 * The warp bitonic network is emulated lane by lane (32 lanes), so ties and the
 * final row order are those of the device code. */
void orc_process_observation_rays(float * obs, u32 * num_rows, u32 max_rows, uint64_t stride, u32 N)
{
    #pragma omp parallel for schedule(static)
    for (int64_t id = 0; id < (int64_t)N; ++id) {
        u32 rows = std::min(num_rows[id], max_rows);
        float ray[32][4]; int key[32]; float theta[32];
        for (int tx = 0; tx < 32; ++tx) {
            key[tx] = tx; theta[tx] = -1.0f;
            ray[tx][0] = ray[tx][1] = ray[tx][2] = ray[tx][3] = 0.0f;
            if ((u32)tx < rows) {
                std::memcpy(ray[tx], obs + 4 * (tx * stride + id), 16);
                theta[tx] = (ray[tx][3] >= 0.0f) ? ray[tx][2] : -1.0f;
            }
        }
        for (int i = 0; (1u << i) < rows; ++i) {
            for (int s = 1 << i; s > 0; s >>= 1) {
                float nt[32]; int nk[32];
                for (int tx = 0; tx < 32; ++tx) {
                    bool asc = (tx >> (i + 1)) % 2;
                    bool dir = (tx % (s << 1)) < s;
                    float otheta = theta[tx ^ s]; int okey = key[tx ^ s];
                    nt[tx] = theta[tx]; nk[tx] = key[tx];
                    if (otheta != theta[tx] && ((otheta > theta[tx]) == (asc ^ dir))) { nt[tx] = otheta; nk[tx] = okey; }
                }
                std::memcpy(theta, nt, sizeof(nt)); std::memcpy(key, nk, sizeof(nk));
            }
        }
        u32 num_valid = 0;
        for (int tx = 0; tx < 32; ++tx) num_valid += (theta[tx] >= 0.0f) ? 1 : 0;
        if (num_valid < rows) num_rows[id] = num_valid;
        for (u32 tx = 0; tx < num_valid && tx < 32; ++tx)
            std::memcpy(obs + 4 * (tx * stride + id), ray[key[tx]], 16);
    }
}

/* ---- K3: evaluate_observation_rays, :266-291 ---- */
void orc_evaluate_observation_rays(float const * obs, u32 const * num_rows, u32 max_rows,
    uint64_t stride, u32 N, float const * heur4, float * recons)
{
    Heur h = {heur4[0], heur4[1], heur4[2], heur4[3]};
    #pragma omp parallel for schedule(dynamic, 256)
    for (int64_t id = 0; id < (int64_t)N; ++id) {
        u32 rows = std::min(num_rows[id], max_rows);
        float const * col = obs + 4 * id;
        float recon = rows >= 1 ? 0.0f : -1.0f;
        for (u32 i = 1; i < rows; ++i)
            recon += heuristic(col, stride, i, col + 4 * stride * i, h);
        recons[id] = recon;
    }
}

/* ---- K8: calculate_func_recons, :496-510 ---- */
void orc_calculate_func_recons(float const * recons, float target, float * wrecons, u32 N)
{
    for (u32 i = 0; i < N; ++i) wrecons[i] = func(recons[i], target);
}

/* ---- K9: cacc::reduction::{sum,min,max}, libs/cacc/reduction.h:132-162 ----
 * (float atomics make the device order unspecified; the oracle sums in double
 * and the gate is 1e-5 relative). out = {sum, min, max} */
void orc_reduce(float const * x, u32 N, double * out)
{
    double s = 0.0; float mn = INF, mx = -INF;
    for (u32 i = 0; i < N; ++i) { s += x[i]; mn = std::min(mn, x[i]); mx = std::max(mx, x[i]); }
    out[0] = s; out[1] = mn; out[2] = mx;
}

/* ---- K4: populate_spherical_histogram (6-arg), :293-336 ---- */
void orc_populate_spherical_histogram(void * bvh, void * kd, float const * view_pos3, float max_distance,
    float const * verts, float const * normals, float const * values, float const * qualities, u32 N,
    double * sphere_hist /* accumulated in double; S entries */, u32 * bins /* optional [N], NAI if culled */,
    float * scores /* optional [N] */)
{
    Tree const & tree = *(Tree *)bvh; KDTree const & kdt = *(KDTree *)kd;
    V3 view_pos = {view_pos3[0], view_pos3[1], view_pos3[2]};
    for (u32 id = 0; id < N; ++id) {
        if (bins) bins[id] = NAI;
        if (scores) scores[id] = 0.0f;
        V3 v = ld3(verts, id), n = ld3(normals, id);
        V3 v2c = sub(view_pos, v);
        float l = norm(v2c);
        float scale = 1.0f - (l / max_distance);
        if (scale <= 0.0f) continue;
        V3 v2cn = divs(v2c, l);
        float ctheta = dot(v2cn, n);
        if (ctheta < 0.087f) continue;
        if (!visible(tree, v, v2cn, l, nullptr)) continue;
        float capture_difficulty = std::max(qualities[id], 0.0f);
        float min_theta = std::min(values[id], 1.484f);
        float scaling = (PI / 2.0f) / ((PI / 2.0f) - min_theta);
        float theta = std::acos(std::min(std::max(ctheta, 0.0f), 1.0f));
        float rel_theta = std::max(theta - min_theta, 0.0f) * scaling;
        float score = capture_difficulty * std::cos(rel_theta) * scale;
        u32 idx = NAI;
        kd_find_nn(kdt, neg(v2cn), &idx, nullptr);
        sphere_hist[idx] += score;
        if (bins) bins[id] = idx;
        if (scores) scores[id] = score;
    }
}

/* ---- K5: populate_spherical_histogram (9-arg), :338-390 ---- */
void orc_populate_spherical_histogram_recon(void * bvh, void * kd, float const * view_pos3,
    float max_distance, float target_recon,
    float const * verts, float const * normals, u32 N,
    float const * obs, u32 const * num_rows, u32 max_rows, uint64_t stride,
    float const * recons, float const * heur4,
    double * sphere_hist, u32 * bins, float * deltas)
{
    Tree const & tree = *(Tree *)bvh; KDTree const & kdt = *(KDTree *)kd;
    Heur h = {heur4[0], heur4[1], heur4[2], heur4[3]};
    V3 view_pos = {view_pos3[0], view_pos3[1], view_pos3[2]};
    for (u32 id = 0; id < N; ++id) {
        if (bins) bins[id] = NAI;
        if (deltas) deltas[id] = 0.0f;
        V3 v = ld3(verts, id), n = ld3(normals, id);
        V3 v2c = sub(view_pos, v);
        float l = norm(v2c);
        float scale = 1.0f - (l / max_distance);
        if (scale <= 0.0f) continue;
        V3 v2cn = divs(v2c, l);
        float ctheta = dot(v2cn, n);
        if (ctheta < 0.087f) continue;
        if (!visible(tree, v, v2cn, l, nullptr)) continue;
        u32 rows = num_rows[id];
        if (rows >= max_rows) continue;
        float recon = recons[id];
        float new_recon = 0.0f;
        if (rows >= 1) {
            V3 rel = relative_direction(v2cn, n);
            float rel4[4] = {rel.x, rel.y, rel.z, scale};
            new_recon = recon + heuristic(obs + 4 * id, stride, rows, rel4, h);
        }
        float delta = func(new_recon, target_recon) - func(recon, target_recon);
        u32 idx = NAI;
        kd_find_nn(kdt, neg(v2cn), &idx, nullptr);
        sphere_hist[idx] += delta;
        if (bins) bins[id] = idx;
        if (deltas) deltas[id] = delta;
    }
}

/* ---- K6: evaluate_spherical_histogram, :392-446 ----
 * hist is [hist_h][hist_w] row-major (pitch == hist_w). */
void orc_evaluate_spherical_histogram(float const * calib9, int width, int height,
    float const * sphere_verts, u32 S, float const * sphere_hist, float * hist, u32 hist_w, u32 hist_h)
{
    for (u32 y = 0; y < hist_h; ++y) for (u32 x = 0; x < hist_w; ++x) {
        float phi = (x / (float)hist_w) * 2.0f * PI;
        float theta = (0.5f + (y / (float)hist_h) / 2.0f) * PI;
        float stheta = std::sin(theta);
        V3 view_dir = {stheta * std::cos(phi), stheta * std::sin(phi), std::cos(theta)};
        V3 rz = neg(view_dir);
        V3 up = {0.0f, 0.0f, 1.0f};
        bool stable = std::abs(dot(up, rz)) < 0.99f;
        if (!stable) up = V3{std::cos(phi), std::sin(phi), 0.0f};
        V3 rx = normalized(cross(up, rz));
        V3 ry = normalized(cross(rz, rx));
        float sum = 0.0f;
        for (u32 i = 0; i < S; ++i) {
            V3 dir = normalized(ld3(sphere_verts, i));
            if (dot(view_dir, dir) < 0.0f) continue;
            V3 v = {dot(rx, dir), dot(ry, dir), dot(rz, dir)};
            float px, py;
            project(v, calib9, &px, &py);
            if (px < 0.0f || width <= px || py < 0.0f || height <= py) continue;
            sum += sphere_hist[i];
        }
        hist[(std::size_t)y * hist_w + x] = sum;
    }
}

/* ---- K7: estimate_capture_difficulty, :448-494 ---- */
void orc_estimate_capture_difficulty(void * bvh, float max_distance, u32 mesh_size,
    float const * sphere_verts, u32 S, float const * verts, float const * normals, u32 N,
    float * values, float * qualities)
{
    Tree const & tree = *(Tree *)bvh;
    #pragma omp parallel for schedule(dynamic, 16)
    for (int64_t id = 0; id < (int64_t)N; ++id) {
        V3 v = ld3(verts, id), n = ld3(normals, id);
        float l = max_distance;
        float sum = 0.0f, max_ctheta = 0.0f;
        for (u32 i = 0; i < S; ++i) {
            V3 dir = normalized(ld3(sphere_verts, i));
            float ctheta = dot(dir, n);
            if (ctheta < 0.087f) continue;
            Ray ray; ray.o = v; ray.d = dir; ray.tmin = l * 0.001f; ray.tmax = l;
            u32 face = NAI;
            bool hit = traverse(tree, ray, false, &face, nullptr, nullptr);
            if (!hit || face >= mesh_size) { if (ctheta > max_ctheta) max_ctheta = ctheta; }
            else sum += ctheta;
        }
        float mx = (S / 2.0f) * 0.5f;
        float min_theta = std::acos(std::min(std::max(max_ctheta, 0.0f), 1.0f));
        values[id] = min_theta;
        qualities[id] = sum / mx;
    }
}

/* ---- proxy-cloud sampler: apps/generate_proxy_cloud/generate_proxy_cloud.cpp:84-184 ----
 * std::mt19937 / std::uniform_real_distribution<float> are the library's own, as in the reference.
 * math::geom::triangle_area is MVE's (not under /root/reference): |(b - a) x (c - a)| / 2.
 * Returns the number of samples; writes them (Z-order sorted, :144-184) when capacity suffices. */
typedef uint64_t u64;
static inline u64 z_spread(u64 x, u64 mask) {          /* libs/acc/math.h:19-45 */
    x &= mask;
    x = (x * 0x100000001ull) & 0x003F00000000FFFFull;
    x = (x * 0x10001ull) & 0x003F0000FF0000FFull;
    x = (x * 0x101ull) & 0x300F00F00F00F00Full;
    x = (x * 0x11ull) & 0x30C30C30C30C30C3ull;
    return (x * 0x5ull) & 0x9249249249249249ull;
}

u32 orc_sample_proxy_cloud(float const * verts, u32 const * faces, u32 num_faces, float samples,
    float * out_v, float * out_n, u32 capacity)
{
    std::uniform_real_distribution<float> dist(0.0f, 1.0f);
    std::mt19937 gen;
    std::vector<V3> pts, nrm;
    for (std::size_t i = 0; i < 3 * (std::size_t)num_faces; i += 3) {
        gen.seed(i);                                     /* :107 */
        V3 v0 = ld3(verts, faces[i]), v1 = ld3(verts, faces[i + 1]), v2 = ld3(verts, faces[i + 2]);
        V3 c = cross(sub(v1, v0), sub(v2, v0));
        float len = norm(c);
        V3 normal = divs(c, len);
        float area = len / 2.0f;
        float face_samples = area * samples;
        unsigned num_face_samples = face_samples;
        if (dist(gen) < (face_samples - static_cast<float>(num_face_samples))) num_face_samples += 1;
        for (unsigned j = 0; j < num_face_samples; ++j) {
            float r1 = dist(gen);
            float r2 = dist(gen);
            float tmp = std::sqrt(r1);
            float u = 1.0f - tmp;
            float v = r2 * tmp;
            float w = 1.0f - v - u;
            pts.push_back(add(add(mul(v0, u), mul(v1, v)), mul(v2, w)));
            nrm.push_back(normal);
        }
    }
    std::size_t n = pts.size();
    if (n == 0 || n > capacity || !out_v || !out_n) return (u32)n;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (std::size_t i = 0; i < n; ++i) {
        float p[3] = {pts[i].x, pts[i].y, pts[i].z};
        for (int j = 0; j < 3; ++j) { mn[j] = std::min(mn[j], p[j]); mx[j] = std::max(mx[j], p[j]); }
    }
    double scale[3], bias[3];
    for (int i = 0; i < 3; ++i) {
        double div = (mx[i] - mn[i]);
        scale[i] = 1.0 / div * (double(1 << 20) - 1.0);
        bias[i] = - mn[i] / div * (double(1 << 20) - 1.0);
    }
    std::vector<u64> zkey(n);
    for (std::size_t i = 0; i < n; ++i) {
        float p[3] = {pts[i].x, pts[i].y, pts[i].z};
        u32 q[3];
        for (int j = 0; j < 3; ++j) q[j] = scale[j] * p[j] + bias[j];
        zkey[i] = z_spread(q[0], 0x3FFFFFull) | (z_spread(q[1], 0x1FFFFFull) << 1) | (z_spread(q[2], 0x1FFFFFull) << 2);
    }
    std::vector<std::size_t> idx(n);
    std::iota(idx.begin(), idx.end(), 0);
    std::stable_sort(idx.begin(), idx.end(), [&](std::size_t li, std::size_t ri) -> bool {
        if (zkey[li] != zkey[ri]) return (zkey[li] < zkey[ri]);
        float a[3] = {pts[li].x, pts[li].y, pts[li].z}, b[3] = {pts[ri].x, pts[ri].y, pts[ri].z};
        return std::lexicographical_compare(a, a + 3, b, b + 3);
    });
    for (std::size_t i = 0; i < n; ++i) {
        V3 p = pts[idx[i]], q = nrm[idx[i]];
        out_v[3 * i] = p.x; out_v[3 * i + 1] = p.y; out_v[3 * i + 2] = p.z;
        out_n[3 * i] = q.x; out_n[3 * i + 1] = q.y; out_n[3 * i + 2] = q.z;
    }
    return (u32)n;
}

}  // extern "C"
