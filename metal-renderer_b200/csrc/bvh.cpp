// bvh.cpp — binned-SAH BVH2 builder (host, multi-threaded, deterministic output).
//
// Stands in for MPSTriangleAccelerationStructure -rebuild (reference: renderer/Renderer.mm:456-462).  The tree
// is built top-down over triangle bounds with 16-bin SAH on all three axes; subtrees are built in parallel into
// a sparse implicit layout (node i over c triangles: left child at i+1, right child at i+2*leftCount) so that
// no thread ever allocates — the result is identical for any thread count — and then emitted depth-first into
// the 64-byte two-box device nodes.  Boxes are inflated by `pad` so that a float-rounded hit that is a
// mathematical near-miss still reaches its triangle: with that, traversal returns exactly
// min over all triangles of (t, triangle index), independent of the tree (tests check against brute force).
#include "bvh.h"
#include "bvh8.h"
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>

namespace mrt {
namespace {

struct Box {
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; ++a) { lo[a] = FLT_MAX; hi[a] = -FLT_MAX; } }
    void grow(const float* l, const float* h) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], l[a]); hi[a] = std::max(hi[a], h[a]); } }
    float half_area() const { float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2]; return dx * dy + dy * dz + dz * dx; }
};

struct BuildNode { Box box; uint32_t first, count, leftCount; };   // leftCount == 0 => leaf

constexpr int kMaxBins = 64;     // SAH bins per axis: Builder::bins, 4..kMaxBins (16 unless the caller says otherwise)
constexpr uint32_t kParallelCutoff = 1u << 14;

// One record per triangle, partitioned IN PLACE as the tree is built, so that every pass over a node's triangles is a
// sequential read (with bounds kept aside and reached through an index array, the five passes per node were random
// gathers: 430 ms of the 1 M-triangle front on 8 cores).  std::partition's swaps depend on the predicate values only, so
// the resulting triangle order is the one the index array used to end up in.
struct TriRef { float lo[3], hi[3]; uint32_t t; uint32_t pad; };

struct Builder {
    std::vector<TriRef> refs;                    // per-triangle bounds + triangle index, in tree order when the build ends
    std::vector<uint32_t> order;                 // refs[k].t, filled by build_front after the build (the emitters read it)
    std::vector<BuildNode> sparse;
    uint32_t maxLeaf;
    int bins = 16;
    // task pool
    struct Task { uint32_t node, first, count; };
    std::mutex mu; std::condition_variable cv; std::deque<Task> queue; size_t pending = 0; bool parallel = false;

    void push(Task t) { { std::lock_guard<std::mutex> g(mu); queue.push_back(t); ++pending; } cv.notify_one(); }

    void worker() {
        for (;;) {
            Task t;
            {
                std::unique_lock<std::mutex> g(mu);
                cv.wait(g, [&] { return !queue.empty() || pending == 0; });
                if (queue.empty()) return;
                t = queue.front(); queue.pop_front();
            }
            build(t.node, t.first, t.count);
            { std::lock_guard<std::mutex> g(mu); if (--pending == 0) cv.notify_all(); }
        }
    }

    void build(uint32_t ni, uint32_t first, uint32_t count) {
        const int kBins = bins;
        for (;;) {   // iterate on the left child, recurse/spawn the right one
            BuildNode& n = sparse[ni];
            n.first = first; n.count = count; n.leftCount = 0;
            Box b, cb; b.reset(); cb.reset();
            const TriRef* ref = refs.data();
            for (uint32_t k = first; k < first + count; ++k) {
                const float* l = ref[k].lo; const float* h = ref[k].hi;
                b.grow(l, h);
                float c[3] = {0.5f * (l[0] + h[0]), 0.5f * (l[1] + h[1]), 0.5f * (l[2] + h[2])};
                cb.grow(c, c);
            }
            n.box = b;
            if (count == 1) return;
            // binned SAH on each axis
            int bestAxis = -1, bestBin = 0; float bestCost = FLT_MAX;
            for (int a = 0; a < 3; ++a) {
                float ext = cb.hi[a] - cb.lo[a];
                if (!(ext > 0.0f)) continue;
                float scale = kBins / ext;
                Box bins[kMaxBins]; uint32_t cnt[kMaxBins];
                for (int i = 0; i < kBins; ++i) { bins[i].reset(); cnt[i] = 0; }
                for (uint32_t k = first; k < first + count; ++k) {
                    const float* l = ref[k].lo; const float* h = ref[k].hi;
                    int bi = std::min(kBins - 1, int((0.5f * (l[a] + h[a]) - cb.lo[a]) * scale));
                    bins[bi].grow(l, h); ++cnt[bi];
                }
                float rightArea[kMaxBins]; uint32_t rightCnt[kMaxBins];
                Box acc; acc.reset(); uint32_t c = 0;
                for (int i = kBins - 1; i > 0; --i) { if (cnt[i]) acc.grow(bins[i].lo, bins[i].hi); c += cnt[i]; rightArea[i] = c ? acc.half_area() : 0.0f; rightCnt[i] = c; }
                acc.reset(); c = 0;
                for (int i = 0; i < kBins - 1; ++i) {
                    if (cnt[i]) acc.grow(bins[i].lo, bins[i].hi);
                    c += cnt[i];
                    if (!c || !rightCnt[i + 1]) continue;
                    float cost = acc.half_area() * c + rightArea[i + 1] * rightCnt[i + 1];
                    if (cost < bestCost) { bestCost = cost; bestAxis = a; bestBin = i; }
                }
            }
            // leaf decision: SAH with unit node/triangle costs, forced split above maxLeaf
            if (count <= maxLeaf) {
                float parent = b.half_area();
                if (bestAxis < 0 || !(parent > 0.0f) || bestCost / parent + 1.0f >= float(count)) return;
            }
            uint32_t mid;
            if (bestAxis < 0) {
                mid = first + count / 2;           // identical centroids: split the list in half
            } else {
                float scale = kBins / (cb.hi[bestAxis] - cb.lo[bestAxis]); float lo = cb.lo[bestAxis];
                auto it = std::partition(refs.begin() + first, refs.begin() + first + count, [&](const TriRef& r) {
                    return std::min(kBins - 1, int((0.5f * (r.lo[bestAxis] + r.hi[bestAxis]) - lo) * scale)) <= bestBin;
                });
                mid = uint32_t(it - refs.begin());
                if (mid == first || mid == first + count) mid = first + count / 2;
            }
            uint32_t leftCount = mid - first, rightCount = count - leftCount;
            sparse[ni].leftCount = leftCount;
            uint32_t rightNode = ni + 2 * leftCount;
            if (parallel && rightCount >= kParallelCutoff) push(Task{rightNode, mid, rightCount});
            else build(rightNode, mid, rightCount);
            ni = ni + 1; count = leftCount;        // continue with the left child
        }
    }
};

}  // namespace

// Triangle bounds, conservative padding and the binned-SAH binary tree (sparse implicit layout) shared by the
// BVH2 emitter below and the 8-wide collapse of build_bvh8.
struct Front {
    std::vector<float> tlo, thi;
    Builder b;
    float pad = 0.0f;
};
static void build_front(const mrt_vertex* vertices, const uint32_t* indices, uint32_t n, uint32_t maxLeaf, uint32_t threads, Front& f,
                        int bins = 16) {
    if (threads == 0) threads = std::max(1u, std::thread::hardware_concurrency());
    f.tlo.resize(3 * size_t(n)); f.thi.resize(3 * size_t(n));
    std::vector<float>& tlo = f.tlo; std::vector<float>& thi = f.thi;
    Box scene; scene.reset();
    for (uint32_t t = 0; t < n; ++t) {
        const float* p0 = vertices[indices[3 * size_t(t)]].v; const float* p1 = vertices[indices[3 * size_t(t) + 1]].v;
        const float* p2 = vertices[indices[3 * size_t(t) + 2]].v;
        for (int a = 0; a < 3; ++a) {
            tlo[3 * size_t(t) + a] = std::min(p0[a], std::min(p1[a], p2[a]));
            thi[3 * size_t(t) + a] = std::max(p0[a], std::max(p1[a], p2[a]));
        }
        scene.grow(&tlo[3 * size_t(t)], &thi[3 * size_t(t)]);
    }
    float ext = 0.0f, mag = 0.0f;
    for (int a = 0; a < 3 && n; ++a) { ext = std::max(ext, scene.hi[a] - scene.lo[a]); mag = std::max(mag, std::max(std::fabs(scene.lo[a]), std::fabs(scene.hi[a]))); }
    f.pad = 1e-5f * std::max(ext, mag) + 1e-30f;
    if (n == 0) return;
    Builder& b = f.b;
    b.maxLeaf = maxLeaf; b.bins = bins < 4 ? 4 : (bins > kMaxBins ? kMaxBins : bins);
    b.refs.resize(n);
    for (uint32_t t = 0; t < n; ++t) {
        TriRef& r = b.refs[t];
        for (int a = 0; a < 3; ++a) { r.lo[a] = tlo[3 * size_t(t) + a]; r.hi[a] = thi[3 * size_t(t) + a]; }
        r.t = t; r.pad = 0;
    }
    b.sparse.resize(2 * size_t(n));
    b.parallel = threads > 1 && n >= 2 * kParallelCutoff;
    if (b.parallel) {
        b.push(Builder::Task{0, 0, n});
        std::vector<std::thread> pool;
        for (uint32_t t = 0; t < threads; ++t) pool.emplace_back([&] { b.worker(); });
        for (auto& th : pool) th.join();
    } else {
        b.build(0, 0, n);
    }
    b.order.resize(n);
    for (uint32_t k = 0; k < n; ++k) b.order[k] = b.refs[k].t;
    std::vector<TriRef>().swap(b.refs);
}

void build_bvh(const mrt_vertex* vertices, const uint32_t* indices, uint32_t n, uint32_t maxLeaf, uint32_t threads,
               BvhResult& out) {
    auto t0 = std::chrono::steady_clock::now();
    out = BvhResult();
    if (maxLeaf < 1) maxLeaf = 1;
    if (maxLeaf > 7) maxLeaf = 7;
    Front front;
    build_front(vertices, indices, n, maxLeaf, threads, front);
    Builder& b = front.b;
    out.pad = front.pad;

    // device triangle records are written in leaf order below; an empty scene gets one node with two empty leaves
    auto empty_leaf = [](BvhNode& nd, int c) {
        float* q = nd.q;
        if (c == 0) { q[0] = FLT_MAX; q[1] = -FLT_MAX; q[2] = FLT_MAX; q[3] = -FLT_MAX; q[8] = FLT_MAX; q[9] = -FLT_MAX; }
        else        { q[4] = FLT_MAX; q[5] = -FLT_MAX; q[6] = FLT_MAX; q[7] = -FLT_MAX; q[10] = FLT_MAX; q[11] = -FLT_MAX; }
        nd.child[c] = ~0;   // leaf, first 0, count 0
    };
    if (n == 0) {
        BvhNode root{}; empty_leaf(root, 0); empty_leaf(root, 1); out.nodes.push_back(root);
        out.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        return;
    }

    // depth-first emission into device nodes; leaves become triangle runs
    out.tris.resize(n);
    out.nodes.reserve(n);
    const float pad = out.pad;
    const float rootArea = std::max(b.sparse[0].box.half_area(), 1e-30f);
    double sah = 0.0;
    auto write_box = [&](BvhNode& nd, int c, const Box& bx) {
        float* q = nd.q;
        if (c == 0) { q[0] = bx.lo[0] - pad; q[1] = bx.hi[0] + pad; q[2] = bx.lo[1] - pad; q[3] = bx.hi[1] + pad; q[8] = bx.lo[2] - pad; q[9] = bx.hi[2] + pad; }
        else        { q[4] = bx.lo[0] - pad; q[5] = bx.hi[0] + pad; q[6] = bx.lo[1] - pad; q[7] = bx.hi[1] + pad; q[10] = bx.lo[2] - pad; q[11] = bx.hi[2] + pad; }
    };
    auto emit_tris = [&](const BuildNode& leaf) {
        for (uint32_t k = leaf.first; k < leaf.first + leaf.count; ++k) {
            uint32_t t = b.order[k];
            const float* V0 = vertices[indices[3 * size_t(t)]].v; const float* V1 = vertices[indices[3 * size_t(t) + 1]].v;
            const float* V2 = vertices[indices[3 * size_t(t) + 2]].v;
            BvhTri& r = out.tris[k];
            for (int a = 0; a < 3; ++a) { r.v2[a] = V2[a]; r.e1[a] = V0[a] - V2[a]; r.e2[a] = V1[a] - V2[a]; }
            r.index = t; r.pad1 = 0.0f; r.pad2 = 0.0f;
        }
        sah += double(leaf.count) * leaf.box.half_area() / rootArea;
    };
    struct Item { uint32_t sparse, parentOut; int slot; uint32_t depth; };
    std::vector<Item> stack;
    if (b.sparse[0].leftCount == 0) {              // whole scene is one leaf: root with one real + one empty child
        BvhNode root{}; write_box(root, 0, b.sparse[0].box); empty_leaf(root, 1);
        root.child[0] = ~int32_t((b.sparse[0].first << 3) | b.sparse[0].count);
        emit_tris(b.sparse[0]);
        out.nodes.push_back(root); out.depth = 1;
    } else {
        out.nodes.push_back(BvhNode{});
        stack.push_back(Item{0, 0, -1, 1});
        while (!stack.empty()) {
            Item it = stack.back(); stack.pop_back();
            const BuildNode& sn = b.sparse[it.sparse];
            uint32_t self = it.slot < 0 ? 0u : uint32_t(out.nodes.size());
            if (it.slot >= 0) { out.nodes.push_back(BvhNode{}); out.nodes[it.parentOut].child[it.slot] = int32_t(self); }
            out.depth = std::max(out.depth, it.depth);
            sah += sn.box.half_area() / rootArea;
            uint32_t kids[2] = {it.sparse + 1, it.sparse + 2 * sn.leftCount};
            for (int c = 1; c >= 0; --c) {         // push right first so the left subtree is emitted adjacent
                const BuildNode& ch = b.sparse[kids[c]];
                write_box(out.nodes[self], c, ch.box);
                if (ch.leftCount == 0) { out.nodes[self].child[c] = ~int32_t((ch.first << 3) | ch.count); emit_tris(ch); }
                else stack.push_back(Item{kids[c], self, c, it.depth + 1});
            }
        }
    }
    out.sah_cost = float(sah);
    out.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}


// ---- 8-wide collapse (bvh8.h) ------------------------------------------------------------------------------------
// Dynamic programme of Ylitie et al. 2017, section 3.1, over the binary tree: C(n, i) = cheapest way to present
// subtree n as at most i children of a wide node; a child is either a leaf (<= 3 triangles, cost area * count * 0.3)
// or another wide node (cost area * 1 + its 8 children).  Leaves of the binary tree are kept whole.
namespace {

constexpr float kCostNode = 1.0f;
// cost of one triangle test relative to one node test (8 boxes); MRT_BVH8_CPRIM overrides it for A/B sweeps
// Measured (pass U, profiles/r1_u_sweep_leaf_cost.log): with vote scheduling a triangle test runs with ~8 of 32 lanes and a
// node test with ~22, so per useful lane a triangle test costs MORE than a node test; cost 1.0 on a front split down to
// single triangles gives 4.3 triangle + 13.0 node tests per ray instead of 6.3 + 12.1 and +4.5 % rays/s (0.3 was the
// instruction-count ratio).
static const float kCostPrim = getenv("MRT_BVH8_CPRIM") ? (float)atof(getenv("MRT_BVH8_CPRIM")) : 1.0f;
// largest leaf the collapse may form (Bvh8Node::leafTris: three bits per slot); MRT_BVH8_MAXLEAF=1 forces single-triangle leaves
static const uint32_t kMaxLeaf8 = getenv("MRT_BVH8_MAXLEAF") ? (uint32_t)std::min(3, std::max(1, atoi(getenv("MRT_BVH8_MAXLEAF")))) : 3u;

struct Collapse {
    const Builder& b;
    float invRootArea;
    // Tables over the sparse node array, left UNINITIALISED (74 MB at 1 M triangles: the zero fill alone was most of the
    // collapse's time): combine() writes C[n][*] and K[n][0] of every node and K[n][1..7], U[n] of every inner node
    // before anything reads them; leaves never have K[n][1..7] or U[n] read (collect() stops at a leaf first).
    std::unique_ptr<float[]> C;  // [node][i-1], i = 1..7
    std::unique_ptr<uint8_t[]> K;   // [node][0] = 1: single child is a leaf; [node][j-1], j = 2..8: left share of distribute(j)
    std::unique_ptr<uint8_t[]> U;   // bit i: C(n,i) comes from distribute(i) rather than from C(n,i-1)
    explicit Collapse(const Builder& bb) : b(bb) {
        size_t m = bb.sparse.size();
        C.reset(new float[7 * m]); K.reset(new uint8_t[8 * m]); U.reset(new uint8_t[m]);
        invRootArea = 1.0f / std::max(bb.sparse[0].box.half_area(), 1e-30f);
    }
    bool leaf(uint32_t n) const { return b.sparse[n].leftCount == 0; }
    uint32_t left(uint32_t n) const { return n + 1; }
    uint32_t right(uint32_t n) const { return n + 2 * b.sparse[n].leftCount; }

    void solve(uint32_t n) {
        if (!leaf(n)) { solve(left(n)); solve(right(n)); }
        combine(n);
    }
    // The same over the subtrees in parallel: a node's table depends on its two children's only, so the result does
    // not depend on how the recursion is cut into threads.  Subtrees below minCount triangles run serially.
    void solve_parallel(uint32_t n, uint32_t minCount) {
        if (leaf(n) || b.sparse[n].count < minCount) { solve(n); return; }
        std::thread other([&] { solve_parallel(left(n), minCount); });
        solve_parallel(right(n), minCount);
        other.join();
        combine(n);
    }
    void combine(uint32_t n) {
        const BuildNode& nd = b.sparse[n];
        const float A = nd.box.half_area() * invRootArea;
        float* c = &C[7 * size_t(n)]; uint8_t* k = &K[8 * size_t(n)];
        if (leaf(n)) {
            for (int i = 0; i < 7; ++i) c[i] = A * float(nd.count) * kCostPrim;
            k[0] = 1;
            return;
        }
        const uint32_t l = left(n), r = right(n);
        const float* cl = &C[7 * size_t(l)]; const float* cr = &C[7 * size_t(r)];
        float D[9];
        for (int j = 2; j <= 8; ++j) {
            float best = FLT_MAX; int bk = 1;
            for (int kk = 1; kk < j; ++kk) {
                if (kk > 7 || j - kk > 7) continue;
                float v = cl[kk - 1] + cr[j - kk - 1];
                if (v < best) { best = v; bk = kk; }
            }
            D[j] = best; k[j - 1] = uint8_t(bk);
        }
        const float cLeaf = nd.count <= kMaxLeaf8 ? A * float(nd.count) * kCostPrim : FLT_MAX;
        const float cInner = D[8] + A * kCostNode;
        c[0] = std::min(cLeaf, cInner); k[0] = cLeaf <= cInner ? 1 : 0;
        uint8_t u = 0;
        for (int i = 2; i <= 7; ++i) {
            if (D[i] < c[i - 2]) { c[i - 1] = D[i]; u |= uint8_t(1u << i); } else c[i - 1] = c[i - 2];
        }
        U[n] = u;
    }
    bool is_leaf_child(uint32_t n) const { return leaf(n) || K[8 * size_t(n)] == 1; }

    // binary-tree nodes that become the (<= budget) children of a wide node
    void collect(uint32_t m, int budget, uint32_t* kids, int& count) const {
        if (leaf(m) || budget == 1) { kids[count++] = m; return; }
        while (budget >= 2 && !(U[m] & (1u << budget))) --budget;
        if (budget == 1) { kids[count++] = m; return; }
        int kk = K[8 * size_t(m) + budget - 1];
        collect(left(m), kk, kids, count);
        collect(right(m), budget - kk, kids, count);
    }
    void children_of(uint32_t n, uint32_t* kids, int& count) const {      // n is an inner wide node
        count = 0;
        int kk = K[8 * size_t(n) + 7];
        collect(left(n), kk, kids, count);
        collect(right(n), 8 - kk, kids, count);
    }
};

}  // namespace

static void build_bvh8_bins(const mrt_vertex* vertices, const uint32_t* indices, uint32_t n, uint32_t threads, int bins, Bvh8Result& out) {
    auto t0 = std::chrono::steady_clock::now();
    out = Bvh8Result();
    out.bins = bins;
    Front front;
    // The binary front is split down to single triangles (MRT_BVH8_FRONTLEAF=1, default), so that the collapse below is
    // free to choose every leaf size from 1 to kMaxLeaf8 itself; 3 = stop at the binary builder's own SAH leaf rule.
    static const uint32_t frontLeaf = getenv("MRT_BVH8_FRONTLEAF") ? (uint32_t)std::min(3, std::max(1, atoi(getenv("MRT_BVH8_FRONTLEAF")))) : 1u;
    build_front(vertices, indices, n, frontLeaf, threads, front, bins);
    static const bool verbose = getenv("MRT_BVH_VERBOSE") != nullptr;
    auto lap = [&](const char* what) {
        if (verbose) fprintf(stderr, "mrt: wide tree (%d bins): %-8s at %7.1f ms\n", bins, what,
                             std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    };
    lap("front");
    const Builder& b = front.b;
    const float pad = front.pad;
    auto finish = [&] { out.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
    if (n == 0) {                                   // one node without children: every ray misses
        Bvh8Node root{}; root.e[0] = root.e[1] = root.e[2] = 127;
        out.nodes.push_back(root); out.depth = 1; finish();
        return;
    }
    Collapse col(b);
    {
        const uint32_t workers = threads ? threads : std::max(1u, std::thread::hardware_concurrency());
        if (workers > 1 && n >= 4 * kParallelCutoff) col.solve_parallel(0, std::max(kParallelCutoff, n / (4 * workers)));
        else col.solve(0);
    }
    lap("collapse");
    out.sah_cost = col.C[0];
    out.tris.reserve(n);
    out.nodes.reserve(size_t(n) / 4 + 16);

    // ---- emission, level by level (= the order of a breadth-first queue), every level in parallel ----
    // A node's children get consecutive node indices in slot order and its leaf triangles consecutive records in slot
    // order; the bases are running counts over the nodes in level order.  So a level is done in three steps: (1) in
    // parallel, every node's children, padded boxes and slot assignment; (2) a serial prefix sum of inner-child and
    // leaf-triangle counts; (3) in parallel, quantisation, the node record, its triangle records and the next level's
    // items at their pre-assigned places.  The result does not depend on the number of threads.
    if (threads == 0) threads = std::max(1u, std::thread::hardware_concurrency());
    auto parallel_for = [&](size_t count, const std::function<void(size_t, size_t)>& body) {
        const size_t workers = std::min<size_t>(threads, (count + 255) / 256);
        if (workers <= 1) { body(0, count); return; }
        std::vector<std::thread> pool;
        for (size_t w = 0; w < workers; ++w) pool.emplace_back([&, w] { body(count * w / workers, count * (w + 1) / workers); });
        for (auto& th : pool) th.join();
    };
    auto write_tris = [&](const BuildNode& leaf, BvhTri* dst) {
        for (uint32_t k = leaf.first; k < leaf.first + leaf.count; ++k, ++dst) {
            uint32_t t = b.order[k];
            const float* V0 = vertices[indices[3 * size_t(t)]].v; const float* V1 = vertices[indices[3 * size_t(t) + 1]].v;
            const float* V2 = vertices[indices[3 * size_t(t) + 2]].v;
            BvhTri r;
            for (int a = 0; a < 3; ++a) { r.v2[a] = V2[a]; r.e1[a] = V0[a] - V2[a]; r.e2[a] = V1[a] - V2[a]; }
            r.index = t; r.pad1 = 0.0f; r.pad2 = 0.0f;
            *dst = r;
        }
    };

    struct Item { uint32_t bnode; uint32_t outIndex; bool wholeLeaf; };
    struct Plan { uint32_t kids[8]; int nk; int kidAt[8]; Box cb[8]; Box nb; uint32_t inner, tris; uint32_t childBase, triBase; };
    std::vector<Item> level, next;
    std::vector<Plan> plans;
    out.nodes.assign(1, Bvh8Node{});
    out.tris.resize(n);                                  // every triangle is emitted exactly once
    uint32_t triAt = 0;
    level.push_back(Item{0, 0, col.is_leaf_child(0)});
    uint32_t depth = 0;
    while (!level.empty()) {
        ++depth;
        plans.resize(level.size());
        parallel_for(level.size(), [&](size_t lo, size_t hi) {
            for (size_t i = lo; i < hi; ++i) {
                const Item& it = level[i];
                Plan& P = plans[i];
                P.nk = 0;
                if (it.wholeLeaf) { P.kids[0] = it.bnode; P.nk = 1; }          // only the root can be a bare leaf
                else col.children_of(it.bnode, P.kids, P.nk);
                const int nk = P.nk;
                // padded child boxes and their union
                P.nb.reset();
                for (int c = 0; c < nk; ++c) {
                    P.cb[c] = b.sparse[P.kids[c]].box;
                    for (int a = 0; a < 3; ++a) { P.cb[c].lo[a] -= pad; P.cb[c].hi[a] += pad; }
                    P.nb.grow(P.cb[c].lo, P.cb[c].hi);
                }
                // slot assignment: slot s (bit a set = +axis a) should hold the child lying furthest that way, so that
                // visiting slots by descending (s ^ octInv) is roughly front to back for every ray octant
                int slotOf[8]; bool slotUsed[8] = {}, kidDone[8] = {};
                float score[8][8];
                for (int c = 0; c < nk; ++c)
                    for (int s = 0; s < 8; ++s) {
                        float v = 0.0f;
                        for (int a = 0; a < 3; ++a) {
                            float d = 0.5f * (P.cb[c].lo[a] + P.cb[c].hi[a]) - 0.5f * (P.nb.lo[a] + P.nb.hi[a]);
                            v += ((s >> a) & 1) ? d : -d;
                        }
                        score[c][s] = v;
                    }
                for (int r = 0; r < nk; ++r) {
                    int bc = -1, bs = -1; float bv = -FLT_MAX;
                    for (int c = 0; c < nk; ++c) if (!kidDone[c])
                        for (int s = 0; s < 8; ++s) if (!slotUsed[s] && (bc < 0 || score[c][s] > bv)) { bv = score[c][s]; bc = c; bs = s; }
                    slotOf[bc] = bs; kidDone[bc] = true; slotUsed[bs] = true;
                }
                for (int s = 0; s < 8; ++s) P.kidAt[s] = -1;
                for (int c = 0; c < nk; ++c) P.kidAt[slotOf[c]] = c;
                P.inner = 0; P.tris = 0;
                for (int c = 0; c < nk; ++c) {
                    if (col.is_leaf_child(P.kids[c])) P.tris += b.sparse[P.kids[c]].count; else ++P.inner;
                }
            }
        });
        uint32_t nodeAt = uint32_t(out.nodes.size());
        const uint32_t nodeStart = nodeAt;
        for (Plan& P : plans) { P.childBase = nodeAt; P.triBase = triAt; nodeAt += P.inner; triAt += P.tris; }
        out.nodes.resize(nodeAt);
        next.resize(nodeAt - nodeStart);
        parallel_for(level.size(), [&](size_t lo, size_t hi) {
            for (size_t i = lo; i < hi; ++i) {
                const Plan& P = plans[i];
                const int nk = P.nk;
                Bvh8Node nd{};
                for (int a = 0; a < 3; ++a) nd.p[a] = P.nb.lo[a];
                // per-axis power-of-two grid step with 255 * step >= extent; planes rounded outwards
                for (int a = 0; a < 3; ++a) {
                    double ext = double(P.nb.hi[a]) - double(P.nb.lo[a]);
                    int ex = 0;
                    std::frexp(std::max(ext, 1e-37) / 255.0, &ex);
                    for (;; ++ex) {
                        if (ex < -126) ex = -126;
                        const double step = std::ldexp(1.0, ex);
                        bool ok = true;
                        for (int c = 0; c < nk && ok; ++c) {
                            double qh = std::ceil((double(P.cb[c].hi[a]) - double(nd.p[a])) / step);
                            if (qh > 255.0) ok = false;
                        }
                        if (ok || ex >= 127) break;
                    }
                    nd.e[a] = uint8_t(ex + 127);
                    const double step = std::ldexp(1.0, ex);
                    for (int s = 0; s < 8; ++s) {
                        int c = P.kidAt[s];
                        if (c < 0) { nd.qlo[a][s] = 255; nd.qhi[a][s] = 0; continue; }      // empty slot: inverted box
                        double ql = std::floor((double(P.cb[c].lo[a]) - double(nd.p[a])) / step);
                        double qh = std::ceil((double(P.cb[c].hi[a]) - double(nd.p[a])) / step);
                        nd.qlo[a][s] = uint8_t(std::min(255.0, std::max(0.0, ql)));
                        nd.qhi[a][s] = uint8_t(std::min(255.0, std::max(0.0, qh)));
                    }
                }
                // children: inner ones get consecutive node indices in slot order, leaf triangles consecutive records
                nd.childBase = P.childBase;
                nd.triBase = P.triBase;
                uint32_t triOffset = 0, innerRank = 0;
                for (int s = 0; s < 8; ++s) {
                    int c = P.kidAt[s];
                    if (c < 0) continue;
                    const BuildNode& ch = b.sparse[P.kids[c]];
                    if (col.is_leaf_child(P.kids[c])) {
                        nd.leafTris |= ((1u << ch.count) - 1u) << (3 * s);                     // unary count at the slot's three bits
                        write_tris(ch, out.tris.data() + P.triBase + triOffset);
                        triOffset += ch.count;
                    } else {
                        nd.imask |= uint8_t(1u << s);
                        const uint32_t idx = P.childBase + innerRank++;
                        next[idx - nodeStart] = Item{P.kids[c], idx, false};
                    }
                }
                out.nodes[level[i].outIndex] = nd;
            }
        });
        level.swap(next);
        next.clear();
    }
    out.depth = depth;
    lap("emit");
    if (verbose) {                      // FNV-1a over the emitted arrays: the tree must not depend on the thread count
        auto fnv = [](const void* p, size_t bytes) { uint64_t h = 1469598103934665603ull; const unsigned char* c = (const unsigned char*)p;
                                                     for (size_t i = 0; i < bytes; ++i) { h ^= c[i]; h *= 1099511628211ull; } return h; };
        fprintf(stderr, "mrt: wide tree (%d bins): %zu nodes %016llx, %zu triangle records %016llx, depth %u\n", bins, out.nodes.size(),
                (unsigned long long)fnv(out.nodes.data(), out.nodes.size() * sizeof(Bvh8Node)), out.tris.size(),
                (unsigned long long)fnv(out.tris.data(), out.tris.size() * sizeof(BvhTri)), out.depth);
    }
    finish();
}

// ---- host self-test of the wide tree (CPU test-suite; no rendering happens here) ---------------------------------
namespace {
struct HostHit { float t; uint32_t tri; };
inline float hdot(const float* a, const float* b) { return std::fmaf(a[2], b[2], std::fmaf(a[1], b[1], a[0] * b[0])); }
inline void hcross(const float* a, const float* b, float* r) {
    r[0] = std::fmaf(a[1], b[2], -(a[2] * b[1])); r[1] = std::fmaf(a[2], b[0], -(a[0] * b[2])); r[2] = std::fmaf(a[0], b[1], -(a[1] * b[0]));
}
// same operation order as triangle_test in device_trace.cuh
inline void host_triangle_test(const BvhTri& T, const float* o, const float* d, float tmin, float tmax, HostHit& best) {
    float pv[3]; hcross(d, T.e2, pv);
    float det = hdot(T.e1, pv);
    float inv = 1.0f / det;
    float s[3] = {o[0] - T.v2[0], o[1] - T.v2[1], o[2] - T.v2[2]};
    float u = hdot(s, pv) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return;
    float q[3]; hcross(s, T.e1, q);
    float v = hdot(d, q) * inv;
    if (!(v >= 0.0f && u + v <= 1.0f)) return;
    float t = hdot(T.e2, q) * inv;
    if (!(t >= tmin && t <= tmax)) return;
    if (t < best.t || (t == best.t && T.index < best.tri)) { best.t = t; best.tri = T.index; }
}
}  // namespace

// Nearest hit of one ray through the wide tree: the same state machine and node decode as device_trace8.cuh.
static HostHit host_trace8(const Bvh8Result& bvh, const float* o, const float* d, float tmin, float tmax, uint64_t& nv, uint64_t& tt) {
    HostHit best{tmax, 0xFFFFFFFFu};
    const float tiny = 1e-30f;
    Ray8 r;
    r.ox = o[0]; r.oy = o[1]; r.oz = o[2];
    r.invx = 1.0f / (std::fabs(d[0]) < tiny ? std::copysign(tiny, d[0]) : d[0]);
    r.invy = 1.0f / (std::fabs(d[1]) < tiny ? std::copysign(tiny, d[1]) : d[1]);
    r.invz = 1.0f / (std::fabs(d[2]) < tiny ? std::copysign(tiny, d[2]) : d[2]);
    r.negx = std::signbit(d[0]); r.negy = std::signbit(d[1]); r.negz = std::signbit(d[2]);
    r.octInv = 7u - (r.negx | (r.negy << 1) | (r.negz << 2));
    r.magic = 0x47000000u;
    uint32_t gx = 0, gy = 0x01000000u, tx = 0, ty = 0;       // node group (base, slot hits << 24 | imask), triangle group
    uint32_t stack[2 * 64]; int sp = 0;
    for (;;) {
        if (gy & 0xff000000u) {
            uint32_t pr = mrt_perm8(gy >> 24, r.octInv);
            uint32_t slot = uint32_t(31 - __builtin_clz(pr)) ^ r.octInv;
            gy &= ~(0x01000000u << slot);
            if (gy & 0xff000000u) { stack[2 * sp] = gx; stack[2 * sp + 1] = gy; ++sp; }
            uint32_t rel = uint32_t(__builtin_popcount(gy & 0xffu & ((1u << slot) - 1u)));
            const uint32_t* w = reinterpret_cast<const uint32_t*>(&bvh.nodes[gx + rel]);
            ++nv;
            uint32_t hs = bvh8_hit_slots(w, r, tmin, best.t);
            uint32_t imask = w[3] >> 24;
            gx = w[4]; gy = ((hs & imask) << 24) | imask;
            tx = w[5]; ty = bvh8_triangle_mask(w[6], hs & ~imask);
        }
        while (ty) {
            int k = __builtin_ctz(ty); ty &= ty - 1;
            ++tt;
            host_triangle_test(bvh.tris[tx + uint32_t(k)], o, d, tmin, tmax, best);
        }
        if (!(gy & 0xff000000u)) {
            if (sp == 0) break;
            --sp; gx = stack[2 * sp]; gy = stack[2 * sp + 1];
        }
    }
    return best;
}

uint32_t bvh8_selftest(const Bvh8Result& bvh, const float* rays, uint32_t count, uint64_t* nodesVisited, uint64_t* trisTested) {
    uint32_t bad = 0;
    uint64_t nv = 0, tt = 0;
    for (uint32_t i = 0; i < count; ++i) {
        const float* o = rays + 8 * size_t(i); const float* d = o + 4;
        const float tmin = o[3], tmax = d[3];
        HostHit brute{tmax, 0xFFFFFFFFu};
        for (const BvhTri& T : bvh.tris) host_triangle_test(T, o, d, tmin, tmax, brute);
        HostHit best = host_trace8(bvh, o, d, tmin, tmax, nv, tt);
        if (best.tri != brute.tri || !(best.t == brute.t)) ++bad;
    }
    if (nodesVisited) *nodesVisited = nv;
    if (trisTested) *trisTested = tt;
    return bad;
}

// Traversal cost of a tree on a fixed, seeded sample of path-like rays: chains that start at a random point of the scene
// box, travel in a uniformly random direction and continue from wherever they hit (backed off a little), like the
// extension rays of a path (7 of 8 rays of the renderer start on a surface).  Cost = node tests + 1.2 x triangle tests
// per ray — the ratio of warp instructions per useful lane of the two steps of the trace kernel.
static double bvh8_sample_cost(const Bvh8Result& bvh, const float* lo, const float* hi, uint32_t chains, uint32_t length, uint64_t seed) {
    uint64_t state = seed * 0x9E3779B97F4A7C15ull + 0x632BE59BD9B4E019ull;
    auto uni = [&]() {                                 // splitmix64 -> [0, 1)
        uint64_t z = (state += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
        return float(z >> 40) * (1.0f / 16777216.0f);
    };
    uint64_t nv = 0, tt = 0, rays = 0;
    for (uint32_t c = 0; c < chains; ++c) {
        float o[3], d[3];
        for (int a = 0; a < 3; ++a) o[a] = lo[a] + (hi[a] - lo[a]) * uni();
        for (uint32_t k = 0; k < length; ++k) {
            float len;
            do { for (int a = 0; a < 3; ++a) d[a] = uni() * 2.0f - 1.0f; len = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]); } while (len < 1e-3f || len > 1.0f);
            for (int a = 0; a < 3; ++a) d[a] /= len;
            HostHit h = host_trace8(bvh, o, d, 0.0f, INFINITY, nv, tt);
            ++rays;
            if (h.tri == 0xFFFFFFFFu) break;
            const float back = h.t * (1.0f - 1e-4f);
            for (int a = 0; a < 3; ++a) o[a] += d[a] * back;
        }
    }
    return rays ? (double(nv) + 1.2 * double(tt)) / double(rays) : 0.0;
}

// The binned-SAH builder is greedy and its trees differ by several per cent in traversal cost depending on how the
// bins happen to fall on the geometry (pass AP: 16 / 64 bins = 167 / 173 ms of trace time on the 1M-triangle lattice of
// spheres, 213 / 192 ms on the 10M one, 66.5 / 64.2 ms on Cornell-water) — and the SAH cost of the finished tree does
// not predict which way.  So the wide tree is built once per candidate bin count and the cheapest one on the ray
// sample above is kept (sample vs GPU trace time, 64 against 16 bins: -8.5 % / -9.7 % on 10M triangles, -3.8 % / -3.5 % on
// Cornell-water).  Deterministic: same scene, same tree.  MRT_BVH_BINS = "16,64" (default) lists the
// candidates; a single number switches the selection off.
void build_bvh8(const mrt_vertex* vertices, const uint32_t* indices, uint32_t n, uint32_t threads, Bvh8Result& out) {
    auto t0 = std::chrono::steady_clock::now();
    std::vector<int> cand;
    {
        const char* e = getenv("MRT_BVH_BINS");
        std::string spec = e ? e : "16,64";
        size_t at = 0;
        while (at < spec.size()) {
            size_t c = spec.find(',', at);
            if (c == std::string::npos) c = spec.size();
            int v = atoi(spec.substr(at, c - at).c_str());
            if (v >= 4) cand.push_back(v > kMaxBins ? kMaxBins : v);
            at = c + 1;
        }
        if (cand.empty()) cand.push_back(16);
    }
    if (cand.size() == 1 || n < 256) { build_bvh8_bins(vertices, indices, n, threads, cand[0], out); return; }
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t t = 0; t < 3 * n; ++t) {
        const float* v = vertices[indices[t]].v;
        for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], v[a]); hi[a] = std::max(hi[a], v[a]); }
    }
    // the candidates are built and sampled side by side, each with its own worker threads: the serial stretches of one
    // (the top of its front, where one thread bins all the triangles; the prefix sums; the ray sample) run while the
    // other's workers are busy.  Each build is deterministic on its own, so the outcome is the same as one after the other.
    std::vector<Bvh8Result> trees(cand.size());
    std::vector<double> costs(cand.size(), 0.0);
    std::vector<std::thread> builders;
    for (size_t i = 0; i < cand.size(); ++i)
        builders.emplace_back([&, i] {
            build_bvh8_bins(vertices, indices, n, threads, cand[i], trees[i]);
            costs[i] = bvh8_sample_cost(trees[i], lo, hi, 4096, 4, 0x5EEDull);
        });
    for (auto& t : builders) t.join();
    double bestCost = 0.0; size_t best = 0;
    for (size_t i = 0; i < cand.size(); ++i) {
        if (getenv("MRT_BVH_VERBOSE")) fprintf(stderr, "mrt: wide tree with %d SAH bins: %zu nodes, sampled cost %.3f per ray\n", cand[i], trees[i].nodes.size(), costs[i]);
        // the first candidate stays unless another one is at least 2 % cheaper: below that the sample does not decide
        // (1M-triangle lattice: 64 bins 0.5 % cheaper on the sample, 4 % slower on the GPU)
        if (i == 0 || costs[i] < 0.98 * bestCost) { bestCost = costs[i]; best = i; }
    }
    out = std::move(trees[best]);
    out.sample_cost = (float)bestCost;
    out.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace mrt
