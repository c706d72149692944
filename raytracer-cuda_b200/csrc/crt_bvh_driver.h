// crt_bvh_driver.h — orchestration of the BVH build steps over an execution policy.
//
// Exec must provide:
//   template<class T> T* alloc(size_t n);   void release(void* p);
//   template<class F> void for_n(int n, const F& f);          // run f(0..n-1) in parallel
//   void sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t* keys_tmp, uint32_t* vals_tmp, int n,
//                   uint64_t** keys_sorted, uint32_t** vals_sorted);
//   void exclusive_scan(const uint32_t* in, uint32_t* out, int n);
//   void fill_u32(uint32_t* p, uint32_t v, size_t n);
//   void upload(void* dst, const void* src, size_t bytes);    // host -> exec memory
//   void download(void* dst, const void* src, size_t bytes);  // exec memory -> host (synchronises)
//   void ploc_small(const PlocSmallArgs& a);                   // all remaining PLOC rounds of <= CRT_PLOC_SMALL clusters
//   void reserve_tmp(size_t bytes);  template<class T> T* alloc_tmp(size_t n);  void release_tmp(void* p);
//                                                              // temporaries of one build (CudaExec: one arena, so that a
//                                                              // build costs one cudaMalloc + one cudaFree instead of ~30)
// The product uses CudaExec (crt_bvh_build.cu); tests/hostsim uses a sequential HostExec.
#pragma once
#include <cstdio>
#include <cstdlib>

#include "crt_bvh_build.h"

namespace crt {

struct WideBvh {
    WideNode* nodes = nullptr;   // exec memory, num_nodes entries (capacity node_capacity)
    PrimRec* prims = nullptr;    // exec memory, num_prims entries, leaf order
    uint32_t num_nodes = 0;
    uint32_t num_prims = 0;      // all primitive records: the BVH's in leaf order, then the giants
    uint32_t giant_base = 0;     // index of the first giant record (= primitives inside the BVH)
    uint32_t giant_count = 0;    // primitives kept out of the BVH and tested by every ray (CRT_MAX_GIANTS)
    uint32_t depth = 0;          // wide levels
    uint32_t ploc_rounds = 0;
};

// The last rounds of PLOC work on a few hundred clusters each (Cornell + blob: 63 of the 78 rounds have fewer than
// 2 200), where a round costs its launches and its blocking read-back, not its work. Once the cluster count is at most
// CRT_PLOC_SMALL the rest of the rounds run in ONE launch of one CTA (k_ploc_small, crt_bvh_build.cu: the same
// functors, __syncthreads between the steps, a block scan for the compaction).
#ifndef CRT_PLOC_SMALL
#define CRT_PLOC_SMALL 2048
#endif
struct PlocSmallArgs {
    f4* blo;
    f4* bhi;
    uint32_t* bcount;
    uint32_t* cl_in;          // n cluster ids, Morton order
    uint32_t* cl_out;
    uint32_t* nearest;
    uint32_t* merged;
    uint32_t* flags;
    uint32_t* pos;
    uint32_t* node_counter;   // binary nodes allocated so far
    uint32_t* total;          // scratch word of PlocCompactFn
    uint32_t* result;         // [0] rounds run (0xffffffff: a round made no progress), [1] id of the root cluster
    int n;
};

// The rounds of ploc_small as a plain loop over an Exec whose memory the host can read (tests/hostsim's HostExec).
template <class Exec>
void ploc_small_on_host(Exec& ex, const PlocSmallArgs& a) {
    uint32_t* cl_in = a.cl_in;
    uint32_t* cl_out = a.cl_out;
    int n = a.n;
    uint32_t rounds = 0;
    while (n > 1) {
        PlocNearestFn fn;
        fn.blo = a.blo; fn.bhi = a.bhi; fn.cluster = cl_in; fn.nearest = a.nearest; fn.n = n;
        ex.for_n(n, fn);
        PlocMergeFn fm;
        fm.blo = a.blo; fm.bhi = a.bhi; fm.bcount = a.bcount; fm.cluster = cl_in; fm.nearest = a.nearest; fm.merged = a.merged;
        fm.flags = a.flags; fm.node_counter = a.node_counter;
        ex.for_n(n, fm);
        ex.exclusive_scan(a.flags, a.pos, n);
        PlocCompactFn fc;
        fc.merged = a.merged; fc.flags = a.flags; fc.pos = a.pos; fc.cluster_out = cl_out; fc.total = a.total; fc.n = n;
        ex.for_n(n, fc);
        const uint32_t total = *a.total;
        if ((int)total >= n || total == 0) { a.result[0] = 0xffffffffu; a.result[1] = 0u; return; }
        n = (int)total;
        uint32_t* t = cl_in; cl_in = cl_out; cl_out = t;
        rounds++;
    }
    a.result[0] = rounds;
    a.result[1] = cl_in[0];
}

// Temporaries of one build: released on every way out of build_wide_bvh (the early error returns included).
template <class Exec>
struct BuildScratch {
    Exec& ex;
    void* ptr[32];
    int n = 0;
    explicit BuildScratch(Exec& e) : ex(e) {}
    ~BuildScratch() { while (n > 0) ex.release_tmp(ptr[--n]); }
    template <class T> T* alloc(size_t count) {
        T* p = n < 32 ? ex.template alloc_tmp<T>(count) : nullptr;
        if (p) ptr[n++] = p;
        return p;
    }
    BuildScratch(const BuildScratch&) = delete;
    BuildScratch& operator=(const BuildScratch&) = delete;
};

// Returns 0 on success, CRT_E_* otherwise (then `out` holds no allocation). `in` points into exec memory.
template <class Exec>
int build_wide_bvh(Exec& ex, const SceneInput& in, uint32_t leaf_max, WideBvh& out) {
    BuildScratch<Exec> tmp(ex);
    const uint32_t Nall = in.num_triangles + in.num_spheres;
    if (Nall == 0) return CRT_E_INVALID;
    if (Nall > 0x07ffffffu) return CRT_E_LIMIT;   // primitive record indices travel in 27 bits
    if (leaf_max < 1) leaf_max = 1;
    if (leaf_max > 3) leaf_max = 3;
    const uint32_t N_alloc = Nall;
    // every temporary below, summed: 272 bytes per primitive (boxes 32, Morton keys and values 24, binary nodes 72, cluster
    // lists and flags 24, collapse tables 104, wide-node map 8, + 8 spare) and 30 allocations' worth of alignment
    ex.reserve_tmp((size_t)N_alloc * 272 + 30 * 512);

    f4* plo = tmp.template alloc<f4>(N_alloc);
    f4* phi = tmp.template alloc<f4>(N_alloc);
    int32_t* cbounds = tmp.template alloc<int32_t>(6);
    uint64_t* keys = tmp.template alloc<uint64_t>(N_alloc);
    uint64_t* keys_tmp = tmp.template alloc<uint64_t>(N_alloc);
    uint32_t* vals = tmp.template alloc<uint32_t>(N_alloc);
    uint32_t* vals_tmp = tmp.template alloc<uint32_t>(N_alloc);
    const size_t nb = (size_t)2 * N_alloc;   // 2N-1 binary nodes
    f4* blo = tmp.template alloc<f4>(nb);
    f4* bhi = tmp.template alloc<f4>(nb);
    uint32_t* bcount = tmp.template alloc<uint32_t>(nb);
    uint32_t* cl_a = tmp.template alloc<uint32_t>(N_alloc);
    uint32_t* cl_b = tmp.template alloc<uint32_t>(N_alloc);
    uint32_t* nearest = tmp.template alloc<uint32_t>(N_alloc);
    uint32_t* merged = tmp.template alloc<uint32_t>(N_alloc);
    uint32_t* flags = tmp.template alloc<uint32_t>(N_alloc);
    uint32_t* pos = tmp.template alloc<uint32_t>(N_alloc);
    uint32_t* counters = tmp.template alloc<uint32_t>(4);   // [0] wide nodes [1] prims [2] binary nodes [3] ploc total
    if (!plo || !phi || !cbounds || !keys || !keys_tmp || !vals || !vals_tmp || !blo || !bhi || !bcount || !cl_a || !cl_b ||
        !nearest || !merged || !flags || !pos || !counters)
        return CRT_E_NOMEM;

    // 1. boxes
    {
        PrimBoxFn f;
        f.in = in; f.plo = plo; f.phi = phi;
        ex.for_n((int)Nall, f);
    }
    // 1b. giants: repeatedly take out the largest primitive while it is CRT_GIANT_RATIO x larger than the
    //     bounds of everything else
    GiantSet giants;
    giants.count = 0;
    for (int k = 0; k < CRT_MAX_GIANTS; k++) giants.id[k] = 0xffffffffu;
    {
        unsigned long long* key = tmp.template alloc<unsigned long long>(1);
        if (!key) return CRT_E_NOMEM;
        // candidates: the CRT_MAX_GIANTS largest primitives, largest first; accept the longest prefix whose
        // smallest member is still CRT_GIANT_RATIO x larger than the bounds of everything outside the prefix
        GiantSet cand_set = giants, accepted = giants;
        while (cand_set.count < CRT_MAX_GIANTS && Nall - cand_set.count >= 2) {
            const unsigned long long zero = 0ull;
            ex.upload(key, &zero, sizeof(zero));
            GiantScanFn fs;
            fs.plo = plo; fs.phi = phi; fs.skip = cand_set; fs.key = key;
            ex.for_n((int)Nall, fs);
            unsigned long long hk = 0ull;
            ex.download(&hk, key, sizeof(hk));
            const uint32_t cand = (uint32_t)(hk & 0xffffffffull);
            const float cand_extent = bits_f((uint32_t)(hk >> 32));
            {   // insert keeping ids ascending
                uint32_t pos = cand_set.count;
                while (pos > 0 && cand_set.id[pos - 1] > cand) { cand_set.id[pos] = cand_set.id[pos - 1]; pos--; }
                cand_set.id[pos] = cand;
                cand_set.count++;
            }
            const int32_t init[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int32_t)0x80000000, (int32_t)0x80000000, (int32_t)0x80000000};
            ex.upload(cbounds, init, sizeof(init));
            RestBoundsFn fr;
            fr.plo = plo; fr.phi = phi; fr.skip = cand_set; fr.bounds = cbounds;
            ex.for_n((int)Nall, fr);
            int32_t hb[6];
            ex.download(hb, cbounds, sizeof(hb));
            const float rest = fmaxf(ordered_float(hb[3]) - ordered_float(hb[0]),
                                     fmaxf(ordered_float(hb[4]) - ordered_float(hb[1]), ordered_float(hb[5]) - ordered_float(hb[2])));
            if (!(rest > 0.f)) break;                                   // the rest is a point: keep it in a tree
            if (cand_extent > CRT_GIANT_RATIO * rest) accepted = cand_set;
            else if (!(cand_extent > rest)) break;                      // candidates no larger than the rest: no longer prefix can qualify
        }
        giants = accepted;
    }
    const uint32_t N = Nall - giants.count;   // primitives inside the BVH
    const int n0 = (int)N;
    {
        const int32_t init[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int32_t)0x80000000, (int32_t)0x80000000, (int32_t)0x80000000};
        ex.upload(cbounds, init, sizeof(init));
        CentroidBoundsFn f;
        f.plo = plo; f.phi = phi; f.giants = giants; f.cbounds = cbounds;
        ex.for_n(n0, f);
    }
    // 2. Morton order
    uint64_t* keys_sorted = keys;
    uint32_t* vals_sorted = vals;
    {
        MortonFn f;
        f.plo = plo; f.phi = phi; f.cbounds = cbounds; f.giants = giants; f.keys = keys; f.vals = vals;
        ex.for_n(n0, f);
        ex.sort_pairs(keys, vals, keys_tmp, vals_tmp, n0, &keys_sorted, &vals_sorted);
    }
    // 3. PLOC
    {
        InitLeavesFn f;
        f.plo = plo; f.phi = phi; f.sorted_prim = vals_sorted; f.blo = blo; f.bhi = bhi; f.bcount = bcount; f.cluster = cl_a;
        ex.for_n(n0, f);
    }
    {
        const uint32_t init[4] = {1u, 0u, N, 0u};
        ex.upload(counters, init, sizeof(init));
    }
    uint32_t* cl_in = cl_a;
    uint32_t* cl_out = cl_b;
    int n = n0;
    uint32_t rounds = 0;
    uint32_t root = 0;
    bool have_root = false;
    while (n > 1) {
        if (n <= CRT_PLOC_SMALL) {   // the rest of the rounds in one launch
            uint32_t* result = tmp.template alloc<uint32_t>(2);
            if (!result) return CRT_E_NOMEM;
            PlocSmallArgs a;
            a.blo = blo; a.bhi = bhi; a.bcount = bcount; a.cl_in = cl_in; a.cl_out = cl_out; a.nearest = nearest; a.merged = merged;
            a.flags = flags; a.pos = pos; a.node_counter = &counters[2]; a.total = &counters[3]; a.result = result; a.n = n;
            ex.ploc_small(a);
            uint32_t res[2] = {0xffffffffu, 0u};
            ex.download(res, result, sizeof(res));
            if (res[0] == 0xffffffffu) return CRT_E_CUDA;   // no progress: cannot happen (see PlocNearestFn)
            if (getenv("CRT_BUILD_TRACE_ROUNDS")) fprintf(stderr, "ploc: last %u rounds (from %d clusters) in one launch\n", res[0], n);
            rounds += res[0];
            root = res[1];
            have_root = true;
            break;
        }
        PlocNearestFn fn;
        fn.blo = blo; fn.bhi = bhi; fn.cluster = cl_in; fn.nearest = nearest; fn.n = n;
        ex.for_n(n, fn);
        PlocMergeFn fm;
        fm.blo = blo; fm.bhi = bhi; fm.bcount = bcount; fm.cluster = cl_in; fm.nearest = nearest; fm.merged = merged;
        fm.flags = flags; fm.node_counter = &counters[2];
        ex.for_n(n, fm);
        ex.exclusive_scan(flags, pos, n);
        PlocCompactFn fc;
        fc.merged = merged; fc.flags = flags; fc.pos = pos; fc.cluster_out = cl_out; fc.total = &counters[3]; fc.n = n;
        ex.for_n(n, fc);
        uint32_t total = 0;
        ex.download(&total, &counters[3], sizeof(uint32_t));
        if ((int)total >= n || total == 0) return CRT_E_CUDA;   // no progress: cannot happen (see PlocNearestFn)
        if (getenv("CRT_BUILD_TRACE_ROUNDS")) fprintf(stderr, "ploc round %u: %d -> %u clusters\n", rounds, n, total);
        n = (int)total;
        uint32_t* t = cl_in; cl_in = cl_out; cl_out = t;
        rounds++;
    }
    if (!have_root) ex.download(&root, cl_in, sizeof(uint32_t));

    // 4. collapse. Every wide node stands for a distinct binary node (an inner one, or a small subtree that
    // was given its own frame, CollapseFn), of which there are 2N-1; N prim records.
    const uint32_t node_cap = 2 * N;
    auto fail_out = [&](int code) {
        ex.release(out.nodes); ex.release(out.prims);
        out = WideBvh();
        return code;
    };
    out.nodes = ex.template alloc<WideNode>(node_cap);
    out.prims = ex.template alloc<PrimRec>(Nall);
    uint32_t* wide_bnode = tmp.template alloc<uint32_t>(node_cap);
    if (!out.nodes || !out.prims || !wide_bnode) return fail_out(CRT_E_NOMEM);
    ex.upload(wide_bnode, &root, sizeof(uint32_t));
    // 4a. the collapse's dynamic programme, bottom-up over the binary tree (leaves 0..N-1, inner nodes N..2N-2)
    uint32_t* parent = tmp.template alloc<uint32_t>(nb);
    uint32_t* arrived = tmp.template alloc<uint32_t>(nb);
    float* ccost = tmp.template alloc<float>(nb * 8);
    uint8_t* cdec = tmp.template alloc<uint8_t>(nb * 8);
    if (!parent || !arrived || !ccost || !cdec) return fail_out(CRT_E_NOMEM);
    ex.fill_u32(arrived, 0u, nb);
    {
        ParentFn fp;
        fp.blo = blo; fp.bhi = bhi; fp.parent = parent; fp.first_inner = N;
        ex.for_n((int)N - 1, fp);
        CollapseCostFn fc;
        fc.blo = blo; fc.bhi = bhi; fc.bcount = bcount; fc.parent = parent; fc.arrived = arrived; fc.cost = ccost; fc.dec = cdec;
        fc.leaf_max = leaf_max; fc.root = root;
        ex.for_n(n0, fc);
    }
    uint32_t begin = 0, end = 1, depth = 0;
    while (begin < end) {
        CollapseFn f;
        f.blo = blo; f.bhi = bhi; f.bcount = bcount; f.in = in; f.nodes = out.nodes; f.prims = out.prims;
        f.wide_bnode = wide_bnode; f.counters = counters; f.level_begin = begin; f.leaf_max = leaf_max; f.dec = cdec;
        ex.for_n((int)(end - begin), f);
        uint32_t c[2];
        ex.download(c, counters, sizeof(c));
        begin = end;
        end = c[0];
        out.num_prims = c[1];
        depth++;
        if (end > node_cap) return fail_out(CRT_E_LIMIT);   // cannot happen: every wide node stands for a distinct binary node (<= 2N-1)
    }
    out.num_nodes = end;
    out.depth = depth;
    out.ploc_rounds = rounds;

    if (out.num_prims != N) return fail_out(CRT_E_CUDA);
    // giants: records appended after the BVH's
    out.giant_base = N;
    out.giant_count = giants.count;
    if (giants.count) {
        GiantRecordsFn f;
        f.in = in; f.giants = giants; f.prims = out.prims; f.base = N;
        ex.for_n((int)giants.count, f);
    }
    out.num_prims = Nall;
    return CRT_OK;
}

}  // namespace crt
