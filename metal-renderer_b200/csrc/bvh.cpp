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
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <mutex>
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

constexpr int kBins = 16;
constexpr uint32_t kParallelCutoff = 1u << 14;

struct Builder {
    const float* tlo; const float* thi;          // per-triangle bounds, 3 floats each
    std::vector<uint32_t> order;
    std::vector<BuildNode> sparse;
    uint32_t maxLeaf;
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
        for (;;) {   // iterate on the left child, recurse/spawn the right one
            BuildNode& n = sparse[ni];
            n.first = first; n.count = count; n.leftCount = 0;
            Box b, cb; b.reset(); cb.reset();
            for (uint32_t k = first; k < first + count; ++k) {
                const float* l = tlo + 3 * size_t(order[k]); const float* h = thi + 3 * size_t(order[k]);
                b.grow(l, h);
                float c[3] = {0.5f * (l[0] + h[0]), 0.5f * (l[1] + h[1]), 0.5f * (l[2] + h[2])};
                cb.grow(c, c);
            }
            n.box = b;
            if (count == 1) return;
            // 16-bin SAH on each axis
            int bestAxis = -1, bestBin = 0; float bestCost = FLT_MAX;
            for (int a = 0; a < 3; ++a) {
                float ext = cb.hi[a] - cb.lo[a];
                if (!(ext > 0.0f)) continue;
                float scale = kBins / ext;
                Box bins[kBins]; uint32_t cnt[kBins];
                for (int i = 0; i < kBins; ++i) { bins[i].reset(); cnt[i] = 0; }
                for (uint32_t k = first; k < first + count; ++k) {
                    const float* l = tlo + 3 * size_t(order[k]); const float* h = thi + 3 * size_t(order[k]);
                    int bi = std::min(kBins - 1, int((0.5f * (l[a] + h[a]) - cb.lo[a]) * scale));
                    bins[bi].grow(l, h); ++cnt[bi];
                }
                float rightArea[kBins]; uint32_t rightCnt[kBins];
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
                auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](uint32_t t) {
                    const float* l = tlo + 3 * size_t(t); const float* h = thi + 3 * size_t(t);
                    return std::min(kBins - 1, int((0.5f * (l[bestAxis] + h[bestAxis]) - lo) * scale)) <= bestBin;
                });
                mid = uint32_t(it - order.begin());
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

void build_bvh(const mrt_vertex* vertices, const uint32_t* indices, uint32_t n, uint32_t maxLeaf, uint32_t threads,
               BvhResult& out) {
    auto t0 = std::chrono::steady_clock::now();
    out = BvhResult();
    if (maxLeaf < 1) maxLeaf = 1;
    if (maxLeaf > 7) maxLeaf = 7;
    if (threads == 0) threads = std::max(1u, std::thread::hardware_concurrency());
    std::vector<float> tlo(3 * size_t(n)), thi(3 * size_t(n));
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
    out.pad = 1e-5f * std::max(ext, mag) + 1e-30f;

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

    Builder b;
    b.tlo = tlo.data(); b.thi = thi.data(); b.maxLeaf = maxLeaf;
    b.order.resize(n);
    for (uint32_t t = 0; t < n; ++t) b.order[t] = t;
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

}  // namespace mrt
