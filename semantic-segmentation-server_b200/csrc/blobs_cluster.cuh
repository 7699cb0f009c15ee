// K5, fast path: connected regions of one frame by one thread-block cluster (included by blobs.cu inside
// namespace ssb::{anonymous}; see the notes at the top of blobs.cu for what the tables mean).
// ================= one cluster per frame: labelling in shared memory =========================================
constexpr int kStrips = 8;            // CTAs per cluster = horizontal strips of a frame
#ifndef SSB_STRIP_THREADS
#define SSB_STRIP_THREADS 256
#endif
#ifndef SSB_STRIP_TMA
#define SSB_STRIP_TMA 1
#endif
#ifndef SSB_STRIP_WORD_FAST
#define SSB_STRIP_WORD_FAST 1   // label counts: 0 = every word a lane per pixel; 1 = uniform words are counted by one lane each;
                                // 2 = the same with the label bytes of the next 32 words requested one round ahead (slower)
#endif
#ifndef SSB_STRIP_SHARE_SLOW
#define SSB_STRIP_SHARE_SLOW 1  // label counts: the words that take the pixel path are listed and shared by the CTA's warps
#endif
#ifndef SSB_STRIP_STOP
#define SSB_STRIP_STOP 0   // tuning: n > 0 ends the kernel after its n-th phase (cumulative phase timing)
#endif
constexpr int kStripThreads = SSB_STRIP_THREADS;
constexpr int kStripWarps = kStripThreads / 32;

constexpr int kRootListCap = 2048;    // strip-local roots kept in a list (more: the passes over roots scan the run starts instead)

// Dynamic shared memory of a CTA (byte offsets, 16-byte aligned):
//   bits  [rows + 2][S + 2] u32   mask rows y0-1 .. y0+rows as bit rows, zero guard words left and right
//   st    [rows][S] u32           bit l of word (ly, seg): pixel 32*seg + l starts a horizontal run
//   open  [rows][S] u16           x of the first pixel of the run that enters segment (ly, seg) from the left
//   lid   [rows * roi_w] u16      union-find parents, used at run starts only (a run is named by its first pixel);
//                                 before that: staging buffer of the bulk copy of the mask rows (+32 bytes)
//   list  [kRootListCap] u16 + u32 counter   the strip's local roots
//   hk    [kHistRows] u32 tags + [kHistRows][C] u32   the CTA's label counts of up to kHistRows table rows (row mod
//                                 kHistRows, first come first served), added to the global table once at the end:
//                                 thousands of warps adding to the same few global counters (the big background
//                                 region, the common labels) queue up in L2
//   bar   one mbarrier (staging copy)
constexpr int kHistRows = 16;
constexpr int kWordLabels = 4;        // a one-run word with up to this many labels is counted by one lane (label counts)
struct StripLayout {
    int bits, st, open, lid, list, hk, bar, total;
};
inline StripLayout strip_layout(int rows, int S, int roi_w, int C) {
    auto up16 = [](size_t v) { return static_cast<int>((v + 15) / 16 * 16); };
    StripLayout L;
    int at = 0;
    L.bits = at; at += up16(sizeof(unsigned int) * static_cast<size_t>(rows + 2) * (S + 2));
    L.st = at;   at += up16(sizeof(unsigned int) * static_cast<size_t>(rows) * S);
    L.open = at; at += up16(sizeof(unsigned short) * static_cast<size_t>(rows) * S);
    L.lid = at;  at += up16((static_cast<size_t>(rows) * roi_w + 1) / 2 * 4 + 32);
    L.list = at; at += up16(sizeof(unsigned short) * kRootListCap + 16);
    L.hk = at;   at += up16(sizeof(unsigned int) * kHistRows * (static_cast<size_t>(C) + 1));
    L.bar = at;  at += 16;
    L.total = at;
    return L;
}

struct StripParams {
    BlobParams b;
    int rows;        // rows per strip = ceil(roi_h / kStrips)
    int S;           // 32-pixel segments per row = ceil(roi_w / 32)
    FastDiv divS;    // task -> (row, segment); unused when S == 1
    FastDiv divW;    // pixel index -> (row, x) (roi_w >= 2)
    int *edge;       // scratch [B][kStrips][2][roi_w]: frame-wide id of the local root of each pixel of a strip's first / last row
    StripLayout L;
};

// All threads of all CTAs of the cluster.  arrive.release / wait.acquire order this CTA's earlier writes (its
// own threads' through the __syncthreads) before the other CTAs' later reads, global memory included; a
// device-wide __threadfence() in front of it cost 1-2 us per barrier and is not needed for that.
__device__ __forceinline__ void cluster_barrier() {
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// ---- union-find over 16-bit ids in shared memory ----
__device__ __forceinline__ int lfind(const unsigned short *lid, int a) {
    const volatile unsigned short *v = lid;
    int p = v[a];
    while (p != a) {
        a = p;
        p = v[a];
    }
    return a;
}
__device__ __forceinline__ int lfind_halving(unsigned short *lid, int a) {
    volatile unsigned short *v = lid;
    while (true) {
        const int p = v[a];
        if (p == a) return a;
        const int gp = v[p];
        if (gp == p) return p;
        v[a] = static_cast<unsigned short>(gp);
        a = gp;
    }
}
// atomicMin on one 16-bit entry (compare-and-swap on the word that holds it); returns the old entry.
__device__ __forceinline__ int lmin(unsigned short *lid, int idx, int val) {
    unsigned int *word = reinterpret_cast<unsigned int *>(lid) + (idx >> 1);
    const int sh = (idx & 1) * 16;
    unsigned int old = *reinterpret_cast<volatile unsigned int *>(word);
    while (true) {
        const int cur = static_cast<int>((old >> sh) & 0xffffu);
        if (cur <= val) return cur;
        const unsigned int nw = (old & ~(0xffffu << sh)) | (static_cast<unsigned int>(val) << sh);
        const unsigned int seen = atomicCAS(word, old, nw);
        if (seen == old) return cur;
        old = seen;
    }
}
__device__ __forceinline__ void lunite(unsigned short *lid, int a, int b) {
    while (true) {
        a = lfind_halving(lid, a);
        b = lfind_halving(lid, b);
        if (a == b) return;
        if (a < b) {
            const int t = a;
            a = b;
            b = t;
        }
        const int old = lmin(lid, a, b);
        if (old == a) return;
        a = old;
    }
}

// ---- the frame-wide union-find over local roots, in global memory; other CTAs of the cluster write it while
// ---- this one reads, so loads go to L2 (an SM's L1 is not kept coherent with other SMs' stores) ----
__device__ __forceinline__ int gfind(int *par, int a) {
    while (true) {
        const int p = __ldcg(par + a);
        if (p == a) return a;
        const int gp = __ldcg(par + p);
        if (gp == p) return p;
        par[a] = gp;
        a = gp;
    }
}
__device__ __forceinline__ void gunite(int *par, int a, int b) {
    while (true) {
        a = gfind(par, a);
        b = gfind(par, b);
        if (a == b) return;
        if (a < b) {
            const int t = a;
            a = b;
            b = t;
        }
        const int old = atomicMin(&par[a], b);
        if (old == a) return;
        a = old;
    }
}

// Neighbour relations of the 32 pixels of one segment as bit masks (bit l = pixel 32*seg + l), from the bit
// rows of the mask: `cur` is the segment's own word, W / E the same row shifted by one pixel, U / UW / UE the
// row above, D the row below (all zero outside the ROI).
struct SegBits {
    unsigned int live, cur, fg, has_w, has_e, w_same, Wb, Eb;
};
__device__ __forceinline__ SegBits seg_bits(const unsigned int *row, int seg, int S, unsigned int live_last) {
    SegBits b;
    b.cur = row[seg];
    b.Wb = (b.cur << 1) | (row[seg - 1] >> 31);
    b.Eb = (b.cur >> 1) | (row[seg + 1] << 31);
    b.live = seg == S - 1 ? live_last : 0xffffffffu;
    b.has_w = seg == 0 ? b.live & ~1u : b.live;
    b.has_e = seg == S - 1 ? b.live >> 1 : 0xffffffffu;
    b.fg = b.cur & b.live;
    b.w_same = b.has_w & ~(b.cur ^ b.Wb);
    return b;
}
// What the merge pass has to link for row `row` given the row above `up` (see blob_merge_kernel for why the
// other vertical links are implied by the pixel on the left).
struct MergeBits {
    unsigned int vert, nw, ne;
};
__device__ __forceinline__ MergeBits merge_bits(const SegBits &b, const unsigned int *up, int seg) {
    const unsigned int U = up[seg], UW = (U << 1) | (up[seg - 1] >> 31), UE = (U >> 1) | (up[seg + 1] << 31);
    const unsigned int n_same = b.live & ~(b.cur ^ U);
    const unsigned int nw_same = b.has_w & ~(b.cur ^ UW);
    MergeBits m;
    m.vert = n_same & ~(b.w_same & nw_same);
    const unsigned int lone = b.fg & ~n_same;            // foreground under background: diagonals matter
    m.nw = lone & b.has_w & UW & ~b.w_same;
    m.ne = lone & b.has_e & UE & ~(b.has_e & b.Eb);
    return m;
}


// Everything a phase needs to know about the CTA's strip.
struct Strip {
    unsigned int *mb, *st;
    unsigned short *open, *lid, *list;
    unsigned int *n_roots;
    int S, SW, rw, rows_here, y0, base;
    unsigned int live_last;
    FastDiv divS;
    __device__ __forceinline__ const unsigned int *bit_row(int lr) const { return mb + (lr + 1) * SW + 1; }
    __device__ __forceinline__ unsigned int *bit_row(int lr) { return mb + (lr + 1) * SW + 1; }
    __device__ __forceinline__ void split(int t, int &lr, int &seg) const {
        lr = S == 1 ? t : static_cast<int>(fastdiv(static_cast<uint32_t>(t), divS));
        seg = t - lr * S;
    }
    // strip-local index of the first pixel of the run that holds pixel (ly, x)
    __device__ __forceinline__ int run_start(int ly, int x) const {
        const int seg = x >> 5;
        const unsigned int w = st[ly * S + seg] & (0xffffffffu >> (31 - (x & 31)));
        return ly * rw + (w ? (x & ~31) + 31 - __clz(w) : static_cast<int>(open[ly * S + seg]));
    }
    // what the run of pixel (ly, x) holds: its strip-local root once the run starts are flattened, later
    // (phase 7b) the table row of its region
    __device__ __forceinline__ int of_run(int ly, int x) const { return lid[run_start(ly, x)]; }
};

#ifndef SSB_STRIP_MIN_CTAS
#define SSB_STRIP_MIN_CTAS 4   // 64 registers at 256 threads: 4 CTAs per SM, so 64 frames' clusters are resident at once
#endif
__global__ void __launch_bounds__(kStripThreads, SSB_STRIP_MIN_CTAS) blob_cluster_kernel(const StripParams sp) {
    extern __shared__ __align__(16) unsigned char strip_smem[];
    const BlobParams &p = sp.b;
    const int s = blockIdx.x, f = blockIdx.y;            // cluster = the kStrips CTAs along x
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int rows = sp.rows;
    Strip k;
    k.mb = reinterpret_cast<unsigned int *>(strip_smem + sp.L.bits);
    k.st = reinterpret_cast<unsigned int *>(strip_smem + sp.L.st);
    k.open = reinterpret_cast<unsigned short *>(strip_smem + sp.L.open);
    k.lid = reinterpret_cast<unsigned short *>(strip_smem + sp.L.lid);
    k.list = reinterpret_cast<unsigned short *>(strip_smem + sp.L.list);
    k.n_roots = reinterpret_cast<unsigned int *>(strip_smem + sp.L.list + sizeof(unsigned short) * kRootListCap);
    k.S = sp.S; k.SW = sp.S + 2; k.rw = p.roi_w; k.divS = sp.divS;
    k.y0 = s * rows;
    k.rows_here = max(0, min(rows, p.roi_h - k.y0));
    k.base = k.y0 * k.rw;
    k.live_last = (k.rw & 31) ? (1u << (k.rw & 31)) - 1u : 0xffffffffu;
    const int S = k.S, rw = k.rw, rows_here = k.rows_here, y0 = k.y0, base = k.base;
    const int n = p.roi_h * rw;
    int *par = p.root + static_cast<long long>(f) * n;
    int *slotp = p.slot + static_cast<long long>(f) * n;
    int *edge = sp.edge + static_cast<long long>(f) * kStrips * 2 * rw;
    const uint8_t *mask_f = p.mask + static_cast<long long>(f) * p.H * p.W;
    const uint8_t *lab_f = p.labels + static_cast<long long>(f) * p.H * p.W;
    const long long tab = static_cast<long long>(f) * p.max_regions;
    const int tasks = rows_here * S;

    unsigned int *hk_tag = reinterpret_cast<unsigned int *>(strip_smem + sp.L.hk), *hk_cnt = hk_tag + kHistRows;
    for (int i = tid; i < (rows + 2) * k.SW; i += kStripThreads) k.mb[i] = 0u;
    for (int i = tid; i < kHistRows * (p.C + 1); i += kStripThreads) hk_tag[i] = i < kHistRows ? 0xffffffffu : 0u;
    if (tid == 0) k.n_roots[0] = k.n_roots[1] = 0u;                     // [1]: words left to the pixel path (label counts)
    __syncthreads();
    pdl_wait();
    if (s == 0 && tid == 0) p.count[f] = 0;

    // ---- phase 1: mask rows y0-1 .. y0+rows_here as bits ----
    const int r_lo = y0 > 0 ? -1 : 0;                                   // first / one-past-last local row that exists
    const int r_hi = min(rows_here + 1, p.roi_h - y0);
    // With roi_w == W those rows are one contiguous byte range: it is staged in the (not yet used) id array by
    // ONE bulk async copy of its 16-byte aligned middle plus plain loads of the ragged ends, and the bits come
    // from shared memory.  Otherwise every segment is loaded from global memory, four loads in flight per lane.
    const bool staged = SSB_STRIP_TMA && rows_here > 0 && rw == p.W &&
                        static_cast<size_t>(r_hi - r_lo) * rw + 32 <= static_cast<size_t>(rows) * rw * 2;
    if (staged) {
        uint8_t *stg = reinterpret_cast<uint8_t *>(k.lid);
        uint64_t *bar = reinterpret_cast<uint64_t *>(strip_smem + sp.L.bar);
        const uint8_t *src = mask_f + static_cast<long long>(y0 + r_lo) * rw;
        const int total = (r_hi - r_lo) * rw;
        const int phase = static_cast<int>(reinterpret_cast<uintptr_t>(src) & 15);
        const int head = min(total, (16 - phase) & 15);
        const int body = (total - head) & ~15;
        if (tid == 0) {
            mbar_init(bar, 1);
            mbar_init_fence();
            if (body > 0) {
                mbar_arrive_expect_tx(bar, static_cast<uint32_t>(body));
                bulk_copy_g2s_plain(stg + phase + head, src + head, static_cast<uint32_t>(body), bar);
            }
        }
        for (int i = tid; i < head; i += kStripThreads) stg[phase + i] = src[i];
        for (int i = head + body + tid; i < total; i += kStripThreads) stg[phase + i] = src[i];
        __syncthreads();                                                // barrier initialised, ragged ends in place
        if (body > 0) mbar_wait(bar, 0);
        // A THREAD per 32-pixel word: its 32 bytes (at any byte phase) as nine aligned 32-bit words, four pixels
        // at a time: a byte's top bit set iff the byte is not zero, the four top bits gathered by one multiply.
        // (A warp per word with a ballot cost ~23 warp instructions per word, a fifth of the whole kernel.)
        for (int t = tid; t < (r_hi - r_lo) * S; t += kStripThreads) {
            int lr, seg;
            k.split(t, lr, seg);
            const int n_px = min(32, rw - seg * 32);
            const unsigned int at = static_cast<unsigned int>(phase + lr * rw + seg * 32);
            const unsigned int sh = (at & 3u) * 8u;
            const unsigned int *w32 = reinterpret_cast<const unsigned int *>(stg + (at & ~3u));   // the staging area is 16-byte aligned
            const int n_words = (static_cast<int>(at & 3u) + n_px + 3) >> 2;                        // words that hold a live byte
            unsigned int v[9];
#pragma unroll
            for (int j = 0; j < 9; ++j) v[j] = j < n_words ? w32[j] : 0u;
            unsigned int word = 0u;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const unsigned int px4 = __funnelshift_r(v[j], v[j + 1], sh);
                const unsigned int top = (((px4 & 0x7f7f7f7fu) + 0x7f7f7f7fu) | px4) & 0x80808080u;
                word |= (((top >> 7) * 0x01020408u) >> 24) << (4 * j);
            }
            k.bit_row(lr + r_lo)[seg] = n_px == 32 ? word : word & ((1u << n_px) - 1u);
        }
    } else if (rows_here > 0) {
        const int n_t = (r_hi - r_lo) * S;
        for (int t0 = warp; t0 < n_t; t0 += 4 * kStripWarps) {
            uint8_t v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int t = t0 + u * kStripWarps;
                int lr, seg;
                k.split(min(t, n_t - 1), lr, seg);
                const int x = seg * 32 + lane;
                v[u] = (t < n_t && x < rw) ? mask_f[static_cast<long long>(y0 + r_lo + lr) * p.W + x] : 0;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int t = t0 + u * kStripWarps;
                if (t >= n_t) break;                                    // warp-uniform
                int lr, seg;
                k.split(t, lr, seg);
                const unsigned int word = __ballot_sync(0xffffffffu, v[u] != 0);
                if (lane == 0) k.bit_row(lr + r_lo)[seg] = word;
            }
        }
    }
    __syncthreads();
    if (SSB_STRIP_STOP == 1) return;


    // From here to the histograms the work is WORD-level: a thread owns the 32-pixel word (ly, seg) -- all
    // neighbour tests are bit operations on a handful of words -- and loops over the few set bits that need
    // an action (a run start, a link to make).  (A warp per word repeated the same word arithmetic in 32
    // lanes: ~100 M warp instructions per 64 frames, the kernel was issue-bound.)

    // ---- phase 2: runs.  Run starts become union-find nodes (their own parent); every other pixel is found
    // ---- through its run: `st` marks the starts, `open` is the start of the run entering a word ----
    for (int t = tid; t < tasks; t += kStripThreads) {
        int ly, seg;
        k.split(t, ly, seg);
        const SegBits b = seg_bits(k.bit_row(ly), seg, S, k.live_last);
        const unsigned int starts = b.live & ~b.w_same;                // column 0 has no west neighbour: it starts a run
        k.st[t] = starts;
        for (unsigned int q = starts; q; q &= q - 1) {
            const int i = ly * rw + seg * 32 + __ffs(q) - 1;
            k.lid[i] = static_cast<unsigned short>(i);
        }
    }
    __syncthreads();
    if (S <= 32) {
        // a warp per row, a lane per word: the run that enters a word started in the nearest word to the left
        // that holds a start (word 0 always does: column 0)
        for (int ly = warp; ly < rows_here; ly += kStripWarps) {
            const unsigned int mine = lane < S ? k.st[ly * S + lane] : 0u;
            const unsigned int holders = __ballot_sync(0xffffffffu, mine != 0u) & ((1u << lane) - 1u);
            const int from = holders ? 31 - __clz(holders) : 0;
            const int start = __shfl_sync(0xffffffffu, mine ? lane * 32 + 31 - __clz(mine) : 0, from);
            if (lane < S) k.open[ly * S + lane] = static_cast<unsigned short>(lane ? start : 0);
        }
    } else {
        for (int t = tid; t < tasks; t += kStripThreads) {
            int ly, seg;
            k.split(t, ly, seg);
            int sg = seg - 1;
            while (sg > 0 && k.st[ly * S + sg] == 0u) --sg;
            k.open[t] = static_cast<unsigned short>(seg == 0 ? 0 : sg * 32 + 31 - __clz(k.st[ly * S + sg]));
        }
    }
    __syncthreads();

    // ---- phase 3: merge inside the strip: the vertical / diagonal links that the pixel on the left does not imply ----
    for (int t = tid + S; t < tasks; t += kStripThreads) {              // rows 1 ..
        int ly, seg;
        k.split(t, ly, seg);
        const SegBits b = seg_bits(k.bit_row(ly), seg, S, k.live_last);
        const MergeBits m = merge_bits(b, k.bit_row(ly - 1), seg);
        for (unsigned int q = m.vert | m.nw | m.ne; q; q &= q - 1) {
            const int l = __ffs(q) - 1, x = seg * 32 + l;
            const int mine = k.run_start(ly, x);
            if ((m.vert >> l) & 1u) lunite(k.lid, mine, k.run_start(ly - 1, x));
            if ((m.nw >> l) & 1u) lunite(k.lid, mine, k.run_start(ly - 1, x - 1));
            if ((m.ne >> l) & 1u) lunite(k.lid, mine, k.run_start(ly - 1, x + 1));
        }
    }
    __syncthreads();
    if (SSB_STRIP_STOP == 2) return;

    // ---- phase 4: flatten the run starts; strip-local roots enter the frame-wide union-find and the root list;
    // ---- the first and last row publish the frame-wide id of every pixel's local root ----
    for (int t = tid; t < tasks; t += kStripThreads) {
        unsigned int q = k.st[t];
        if (!q) continue;
        int ly, seg;
        k.split(t, ly, seg);
        for (; q; q &= q - 1) {
            const int i = ly * rw + seg * 32 + __ffs(q) - 1;
            const int r = lfind_halving(k.lid, i);
            k.lid[i] = static_cast<unsigned short>(r);  // a concurrent halving step may still re-point it at an older ancestor
            if (r == i) {
                par[base + i] = base + i;
                const unsigned int at = atomicAdd(k.n_roots, 1u);
                if (at < kRootListCap) k.list[at] = static_cast<unsigned short>(i);
            }
        }
    }
    __syncthreads();
    // second pass without any re-pointing in flight: afterwards every run start holds EXACTLY its strip-local root
    for (int t = tid; t < tasks; t += kStripThreads) {
        unsigned int q = k.st[t];
        if (!q) continue;
        int ly, seg;
        k.split(t, ly, seg);
        for (; q; q &= q - 1) {
            const int i = ly * rw + seg * 32 + __ffs(q) - 1;
            k.lid[i] = static_cast<unsigned short>(lfind(k.lid, i));
        }
    }
    __syncthreads();
    if (SSB_STRIP_STOP == 3) return;
    if (rows_here > 0) {
        for (int t = warp; t < 2 * S; t += kStripWarps) {
            const int which = t >= S, seg = t - which * S, ly = which ? rows_here - 1 : 0;
            const int x = seg * 32 + lane;
            if (x < rw) edge[(s * 2 + which) * rw + x] = base + k.of_run(ly, x);
        }
    }
    cluster_barrier();
    if (SSB_STRIP_STOP == 4) return;

    // ---- phase 5: links across the border with the strip above ----
    if (s > 0 && rows_here > 0) {
        const int *up_id = edge + ((s - 1) * 2 + 1) * rw;
        for (int seg = tid; seg < S; seg += kStripThreads) {
            const SegBits b = seg_bits(k.bit_row(0), seg, S, k.live_last);
            const MergeBits m = merge_bits(b, k.bit_row(-1), seg);
            for (unsigned int q = m.vert | m.nw | m.ne; q; q &= q - 1) {
                const int l = __ffs(q) - 1, x = seg * 32 + l;
                const int mine = base + k.of_run(0, x);
                if ((m.vert >> l) & 1u) gunite(par, mine, __ldcg(up_id + x));
                if ((m.nw >> l) & 1u) gunite(par, mine, __ldcg(up_id + x - 1));
                if ((m.ne >> l) & 1u) gunite(par, mine, __ldcg(up_id + x + 1));
            }
        }
    }
    cluster_barrier();
    if (SSB_STRIP_STOP == 5) return;

    // the strip's local roots: from the list, or (more roots than it holds) from the run starts
    const unsigned int n_roots = *k.n_roots;
    auto for_each_root = [&](auto &&fn) {
        if (n_roots <= kRootListCap) {
            for (unsigned int j = tid; j < n_roots; j += kStripThreads) fn(static_cast<int>(k.list[j]));
        } else {
            for (int t = warp; t < tasks; t += kStripWarps) {
                if (!((k.st[t] >> lane) & 1u)) continue;
                int ly, seg;
                k.split(t, ly, seg);
                const int i = ly * rw + seg * 32 + lane;
                if (k.lid[i] == i) fn(i);
            }
        }
    };
    auto row_x = [&](int i, int &ly, int &x) {
        ly = rw == 1 ? i : static_cast<int>(fastdiv(static_cast<uint32_t>(i), sp.divW));
        x = i - ly * rw;
    };

    // ---- phase 6: a table row for every region, handed out by the CTA that holds the region's first pixel ----
    for_each_root([&](int i) {
        if (gfind(par, base + i) != base + i) return;
        const int row = atomicAdd(&p.count[f], 1);
        slotp[base + i] = row;
        if (row >= p.max_regions) return;
        int ly, x;
        row_x(i, ly, x);
        p.first[tab + row] = base + i;
        p.flags[tab + row] = (k.bit_row(ly)[x >> 5] >> (x & 31)) & 1u;
        unsigned int *h = p.hist + (tab + row) * p.C;
        for (int c = 0; c < p.C; ++c) h[c] = 0u;
        if (p.ring) {
            unsigned int *g = p.ring + (tab + row) * p.C;
            for (int c = 0; c < p.C; ++c) g[c] = 0u;
        }
    });
    cluster_barrier();

    // ---- phase 7: every local root learns its region's row; first pixels learn the enclosing region ----
    for_each_root([&](int i) {
        const int g = gfind(par, base + i);
        const int row = __ldcg(slotp + g);
        if (g != base + i) {
            slotp[base + i] = row;
        } else if (row < p.max_regions) {
            int ly, x;
            row_x(i, ly, x);
            // the region that encloses this one owns the pixel left of its first pixel
            p.parent[tab + row] = x > 0 ? gfind(par, base + k.of_run(ly, x - 1)) : -1;
        }
    });
    // The label counts below look at 32 words at a time, a lane per word, and need the words' label bytes (variant 2:
    // those of a warp's first 32 words are requested here, so that they arrive while the cluster waits).
    const int t_begin = static_cast<int>(static_cast<long long>(tasks) * warp / kStripWarps);
    const int t_end = static_cast<int>(static_cast<long long>(tasks) * (warp + 1) / kStripWarps);
    struct WordLook {
        unsigned int v[9];      // the word's 32 label bytes as aligned 32-bit words
        unsigned int off;       // byte phase of the first label in v[0]
        unsigned int n_px;      // live pixels of the word; 0: not a candidate (several runs, ring work, past the range)
        unsigned int starts;
    };
    // A word is a candidate for being counted whole if it lies inside one run and none of its pixels is foreground
    // next to background (no ring work); whether its pixels carry one label is seen when the bytes are there.
    auto look_ahead = [&](int t0) {
        WordLook q;
        q.n_px = 0u; q.off = 0u; q.starts = 0u;
#pragma unroll
        for (int j = 0; j < 9; ++j) q.v[j] = 0u;
        const int tw = t0 + lane;
        if (!SSB_STRIP_WORD_FAST || tw >= t_end) return q;
        int ly, seg;
        k.split(tw, ly, seg);
        const SegBits b = seg_bits(k.bit_row(ly), seg, S, k.live_last);
        q.starts = k.st[tw] & b.live;
        bool cand = (q.starts & ~1u) == 0u;                              // one run from the first live pixel to the last
        if (cand && p.ring && b.fg) {
            const int gy = y0 + ly;
            const unsigned int U = gy > 0 ? k.bit_row(ly - 1)[seg] : 0xffffffffu;
            const unsigned int D = gy + 1 < p.roi_h ? k.bit_row(ly + 1)[seg] : 0xffffffffu;
            cand = (b.fg & ((b.has_w & ~b.Wb) | (b.has_e & ~b.Eb) | ~U | ~D)) == 0u;
        }
        if (!cand) return q;
        q.n_px = __popc(b.live);
        // at least one byte of every 32-bit word loaded is a live pixel of this word
        const uint8_t *src = lab_f + static_cast<long long>(y0 + ly) * p.W + seg * 32;
        q.off = static_cast<unsigned int>(reinterpret_cast<uintptr_t>(src) & 3u);
        const unsigned int *w32 = reinterpret_cast<const unsigned int *>(src - q.off);
        const int n_bytes = static_cast<int>(q.off + q.n_px);
#pragma unroll
        for (int j = 0; j < 9; ++j)
            if (4 * j < n_bytes) q.v[j] = w32[j];
        return q;
    };
    WordLook look = look_ahead(SSB_STRIP_WORD_FAST == 2 ? t_begin : t_end);
    if (p.ring) cluster_barrier(); else __syncthreads();
    if (SSB_STRIP_STOP == 6) return;


    // ---- phase 7b: every run start trades its root for the table row of its region (0xFFFF: not in the table),
    // ---- so that the per-pixel pass reads one shared-memory entry per pixel and nothing else ----
    for (int t = tid; t < tasks; t += kStripThreads) {
        unsigned int q = k.st[t];
        if (!q) continue;
        int ly, seg;
        k.split(t, ly, seg);
        for (; q; q &= q - 1) {
            const int i = ly * rw + seg * 32 + __ffs(q) - 1;
            const int row = slotp[base + k.lid[i]];                    // written by this CTA itself, phases 6 / 7
            k.lid[i] = static_cast<unsigned short>(row < p.max_regions ? row : 0xFFFF);
        }
    }
    __syncthreads();

    // Everything read from global memory from here on was written before the last cluster barrier and no
    // longer changes: plain (L1-cached) loads.

    auto cross_row = [&](int id) {          // a pixel of another strip, by the published id of its local root
        const int row = slotp[id];
        return row < p.max_regions ? row : 0xFFFF;
    };
    // ---- phase 8: label histograms.  A warp owns a contiguous range of words.  Words that are one run with one
    // ---- label and no ring work are counted whole (a lane per WORD, see below); the others a lane per pixel,
    // ---- with the count of ONE (row, label) pair kept in registers from word to word. ----
    {
        unsigned int cache_key = 0xffffffffu, cache_cnt = 0u;
        // n pixels of (row, label): into the CTA's shared-memory counts if the row has (or gets) a place there
        auto add_hist = [&](unsigned int row, unsigned int lab, unsigned int n_px) {
            const unsigned int slot = row & (kHistRows - 1);
            unsigned int tag = *reinterpret_cast<volatile unsigned int *>(hk_tag + slot);
            if (tag == 0xffffffffu) {
                tag = atomicCAS(hk_tag + slot, 0xffffffffu, row);
                if (tag == 0xffffffffu) tag = row;
            }
            if (tag == row) atomicAdd(hk_cnt + slot * p.C + lab, n_px);
            else atomicAdd(&p.hist[(tab + row) * p.C + lab], n_px);
        };
        auto flush_cache = [&]() {
            if (cache_cnt && lane == 0) add_hist(cache_key >> 8, cache_key & 0xffu, cache_cnt);
            cache_cnt = 0u;
        };
        const unsigned int below = 0xffffffffu >> (31 - lane);           // bits of lanes <= lane
        // the root list is free again: it takes the words that are left to the pixel path (when they all fit)
        const bool listed = SSB_STRIP_SHARE_SLOW && tasks <= kRootListCap;
        unsigned int *n_slow = k.n_roots + 1;
        // one word, a lane per pixel
        auto step = [&](int t, int ly, int seg, int lab) {
            const int x = seg * 32 + lane;
            unsigned int row = 0xFFFFu;
            if (x < rw) {
                const unsigned int mine = k.st[t] & below;
                row = k.lid[ly * rw + (mine ? seg * 32 + 31 - __clz(mine) : static_cast<int>(k.open[t]))];
            }
            const bool counted = row != 0xFFFFu && lab < p.C;           // dead lanes: row = 0xFFFF
            const unsigned int key = (row << 8) | static_cast<unsigned int>(lab);
            const bool hit = counted && key == cache_key;
            cache_cnt += __popc(__ballot_sync(0xffffffffu, hit));
            const unsigned int rest = __ballot_sync(0xffffffffu, counted && !hit);
            if (rest) {
                const unsigned int peers = __match_any_sync(0xffffffffu, counted ? key : 0xfffffffeu - lane);
                const int lead = __ffs(rest) - 1;
                if (__shfl_sync(0xffffffffu, peers, lead) == rest && __popc(rest) >= 16) {
                    flush_cache();                    // one new pair covers the rest of the word: it becomes the cached one
                    cache_key = __shfl_sync(0xffffffffu, key, lead);
                    cache_cnt = __popc(rest);
                } else if (counted && !hit && lane == __ffs(peers) - 1) {
                    add_hist(row, lab, __popc(peers));
                }
            }
            // Ring rows.  Hole contours run through the pixels of the ENCLOSING foreground region that are
            // 4-adjacent to the hole (not through islands inside the hole): such a pixel adds its label to the
            // hole's ring row.  Only words with foreground get here.
            const unsigned int cur = p.ring ? k.bit_row(ly)[seg] : 0u;
            if (cur) {
                const int gy = y0 + ly;
                const SegBits b = seg_bits(k.bit_row(ly), seg, S, k.live_last);
                const unsigned int U = gy > 0 ? k.bit_row(ly - 1)[seg] : 0xffffffffu;
                const unsigned int D = gy + 1 < p.roi_h ? k.bit_row(ly + 1)[seg] : 0xffffffffu;
                const unsigned int west_bg = b.has_w & ~b.Wb, east_bg = b.has_e & ~b.Eb;
                if (((b.fg & (west_bg | east_bg | ~U | ~D)) >> lane) & 1u && counted) {
                    // The neighbour's table row comes from its run (rows outside the strip: the published ids);
                    // it is a hole of THIS region if the region left of its first pixel is this pixel's region
                    // (first pixel in column 0: parent -1, the "hole" is the outside).
                    int hole[4];
                    hole[0] = (west_bg >> lane) & 1u ? k.of_run(ly, x - 1) : 0xFFFF;
                    hole[1] = (east_bg >> lane) & 1u ? k.of_run(ly, x + 1) : 0xFFFF;
                    hole[2] = (U >> lane) & 1u ? 0xFFFF : (ly > 0 ? k.of_run(ly - 1, x) : cross_row(edge[((s - 1) * 2 + 1) * rw + x]));
                    hole[3] = (D >> lane) & 1u ? 0xFFFF : (ly + 1 < rows_here ? k.of_run(ly + 1, x) : cross_row(edge[((s + 1) * 2 + 0) * rw + x]));
                    int mine_first = -2;
#pragma unroll
                    for (int d = 0; d < 4; ++d) {
                        if (hole[d] == 0xFFFF) continue;
                        bool dup = false;
#pragma unroll
                        for (int e = 0; e < d; ++e) dup |= hole[e] == hole[d];
                        if (dup) continue;
                        if (mine_first == -2) mine_first = p.first[tab + row];
                        if (p.parent[tab + hole[d]] != mine_first) continue;   // also: this pixel belongs to an island inside the hole
                        atomicAdd(&p.ring[(tab + hole[d]) * p.C + lab], 1u);
                    }
                }
            }
        };
        auto label_of = [&](int ly, int seg) -> int {
            const int x = seg * 32 + lane;
            return x < rw ? lab_f[static_cast<long long>(y0 + ly) * p.W + x] : 0;
        };
#if SSB_STRIP_WORD_FAST
        // 32 words at a time, first a LANE per word: a candidate word whose pixels carry at most kWordLabels labels is
        // counted by its lane alone -- one-label words of the same (row, label) among the 32 by one lane for all.
        // The other words then take the pixel path.
        for (int t0 = t_begin; t0 < t_end; t0 += 32) {
            const int tw = t0 + lane;
            bool slow_word = tw < t_end;
            unsigned int key = 0xfffffffeu - lane;                       // matches nobody
            if (SSB_STRIP_WORD_FAST == 1) look = look_ahead(t0);
            if (look.n_px) {
                // The word's pixels by label, up to kWordLabels labels: the label of the first pixel not yet
                // accounted for, the pixels that carry it (XOR, top bit of every non-zero byte, four bits gathered
                // by one multiply), until none is left.  A word with more labels takes the pixel path.
                unsigned int w[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) w[j] = __funnelshift_r(look.v[j], look.v[j + 1], 8 * look.off);
                int ly, seg;
                k.split(tw, ly, seg);
                const uint8_t *lab_row = lab_f + static_cast<long long>(y0 + ly) * p.W + seg * 32;
                unsigned int left = look.n_px == 32u ? 0xffffffffu : (1u << look.n_px) - 1u;   // pixels not yet accounted for
                unsigned int labs[kWordLabels], cnts[kWordLabels];
#pragma unroll
                for (int r = 0; r < kWordLabels; ++r) {
                    labs[r] = 0u;
                    cnts[r] = 0u;
                    if (left && !(r >= 2 && __popc(left) > 8)) {          // two labels gone and a quarter still open: give up
                        const unsigned int lab_r = r == 0 ? (w[0] & 0xffu) : lab_row[__ffs(left) - 1];
                        const unsigned int rep = lab_r * 0x01010101u;
                        unsigned int other = 0u;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const unsigned int x4 = w[j] ^ rep;
                            const unsigned int top = (((x4 & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x4) & 0x80808080u;
                            other |= (((top >> 7) * 0x01020408u) >> 24) << (4 * j);
                        }
                        const unsigned int same = left & ~other;
                        labs[r] = lab_r;
                        cnts[r] = __popc(same);
                        left &= other;
                    }
                }
                if (left == 0u) {
                    slow_word = false;
                    const unsigned int row = k.lid[ly * rw + ((look.starts & 1u) ? seg * 32 : static_cast<int>(k.open[tw]))];
                    if (row != 0xFFFFu) {
                        if (cnts[0] == 32u) {
                            if (labs[0] < static_cast<unsigned int>(p.C)) key = (row << 8) | labs[0];
                        } else {
#pragma unroll
                            for (int r = 0; r < kWordLabels; ++r)
                                if (cnts[r] && labs[r] < static_cast<unsigned int>(p.C)) add_hist(row, labs[r], cnts[r]);
                        }
                    }
                }
            }
            if (SSB_STRIP_WORD_FAST == 2) look = look_ahead(t0 + 32);
            const unsigned int peers = __match_any_sync(0xffffffffu, key);
            if (key < 0xffffff00u && lane == __ffs(peers) - 1) add_hist(key >> 8, key & 0xffu, 32u * __popc(peers));
            unsigned int slow = __ballot_sync(0xffffffffu, slow_word);
            if (listed) {
                // the CTA's warps share the pixel-path words afterwards (a warp whose range holds many of them would
                // keep the others waiting)
                if (slow) {
                    unsigned int at = lane == 0 ? atomicAdd(n_slow, static_cast<unsigned int>(__popc(slow))) : 0u;
                    at = __shfl_sync(0xffffffffu, at, 0);
                    if (slow_word) k.list[at + __popc(slow & ((1u << lane) - 1u))] = static_cast<unsigned short>(tw);
                }
            } else if (slow) {
                int ly, seg, nly = 0, nseg = 0;
                int wi = __ffs(slow) - 1;
                slow &= slow - 1;
                k.split(t0 + wi, ly, seg);
                int lab = label_of(ly, seg);
                while (true) {
                    const int nwi = slow ? __ffs(slow) - 1 : -1;         // the next word's labels are requested first
                    int nlab = 0;
                    if (nwi >= 0) {
                        slow &= slow - 1;
                        k.split(t0 + nwi, nly, nseg);
                        nlab = label_of(nly, nseg);
                    }
                    step(t0 + wi, ly, seg, lab);
                    if (nwi < 0) break;
                    wi = nwi; ly = nly; seg = nseg; lab = nlab;
                }
            }
        }
        if (listed) {
            __syncthreads();
            const int n_listed = static_cast<int>(*n_slow);
            int i = warp, ly = 0, seg = 0, lab = 0, t = 0;
            if (i < n_listed) {
                t = k.list[i];
                k.split(t, ly, seg);
                lab = label_of(ly, seg);
            }
            while (i < n_listed) {
                const int ni = i + kStripWarps;                          // the next word's labels are requested first
                int nt = 0, nly = 0, nseg = 0, nlab = 0;
                if (ni < n_listed) {
                    nt = k.list[ni];
                    k.split(nt, nly, nseg);
                    nlab = label_of(nly, nseg);
                }
                step(t, ly, seg, lab);
                i = ni; t = nt; ly = nly; seg = nseg; lab = nlab;
            }
        }
#else
        // every word takes the pixel path, the label bytes of the next three words already requested
        {
            int t = t_begin, ly, seg;
            k.split(t < tasks ? t : 0, ly, seg);
            int l_t = t, l_ly = ly, l_seg = seg;
            auto load_label = [&]() -> int {
                const int v = l_t < t_end ? label_of(l_ly, l_seg) : 0;
                ++l_t;
                if (++l_seg == S) { l_seg = 0; ++l_ly; }
                return v;
            };
            auto advance = [&]() { ++t; if (++seg == S) { seg = 0; ++ly; } };
            int la = load_label(), lb = load_label(), lc = load_label();
            while (t < t_end) {
                step(t, ly, seg, la); advance(); la = load_label();
                if (t >= t_end) break;
                step(t, ly, seg, lb); advance(); lb = load_label();
                if (t >= t_end) break;
                step(t, ly, seg, lc); advance(); lc = load_label();
            }
        }
#endif
        flush_cache();
    }
    __syncthreads();
    for (int e = tid; e < kHistRows * p.C; e += kStripThreads) {
        const unsigned int n_px = hk_cnt[e];
        if (n_px) {
            const int slot = e / p.C;
            atomicAdd(&p.hist[(tab + hk_tag[slot]) * p.C + (e - slot * p.C)], n_px);
        }
    }

    // ---- phase 9 (a thread per word again): background regions that touch the ROI border are the outside, not holes ----
    for (int t = tid; t < tasks; t += kStripThreads) {
        int ly, seg;
        k.split(t, ly, seg);
        const int gy = y0 + ly;
        if (gy != 0 && gy != p.roi_h - 1 && seg != 0 && seg != S - 1) continue;
        const SegBits b = seg_bits(k.bit_row(ly), seg, S, k.live_last);
        const unsigned int bg = b.live & ~b.cur;
        // border pixels of the word that are background.  A full border row: one atomic per run (the word's
        // first border pixel, and every run start among the others); otherwise just the pixels in column 0 / W-1
        unsigned int em;
        if (gy == 0 || gy == p.roi_h - 1) {
            em = bg & (k.st[t] | (bg & (0u - bg)));
        } else {
            em = (seg == 0 ? bg & 1u : 0u) | (seg == S - 1 ? bg & (1u << ((rw - 1) & 31)) : 0u);
        }
        for (unsigned int q = em; q; q &= q - 1) {
            const int row = k.of_run(ly, seg * 32 + __ffs(q) - 1);
            if (row != 0xFFFF) atomicOr(&p.flags[tab + row], 2);
        }
    }
}
