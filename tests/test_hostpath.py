"""The native host-side helpers of the producer loop: batch fold of the region tables (no GPU needed) and the
frame upload that sends page-locked frames from where they are and gathers the others with host threads."""
import numpy as np
import pytest

from test_regions import random_label_map, scipy_table


def test_native_fold_equals_the_python_fold():
    import ssb200
    rng = np.random.default_rng(4)
    R, B = 2048, 40
    count = np.zeros(B, np.int32)
    first, parent, flags = (np.full((B, R), -7, np.int32) for _ in range(3))      # garbage beyond count, as on the GPU
    hist, ring = (np.full((B, R, 21), 0x55555555, np.uint32) for _ in range(2))
    want = []
    for b in range(B):
        mask, lab = random_label_map(rng, b)
        n, f, p, fl, h, rg = scipy_table(mask, lab)
        perm = rng.permutation(n)                       # the kernel hands out rows with an atomic counter
        count[b] = n
        first[b, :n], parent[b, :n], flags[b, :n], hist[b, :n], ring[b, :n] = f[perm], p[perm], fl[perm], h[perm], rg[perm]
        want.append(ssb200.fold_region_histograms(n, f, p, fl, h, rg))
    count[3], want[3] = R + 5, None                     # an overflowed table
    tables = {"count": count, "first": first, "parent": parent, "flags": flags, "hist": hist, "ring": ring}
    scratch = {}
    for threads in (1, 5):                              # the second call reuses the work arrays
        got = ssb200.fold_regions_batch(tables, threads=threads, scratch=scratch)
        for b in range(B):
            if want[b] is None:
                assert got[b] is None
                continue
            assert set(got[b]) == set(want[b]) and set(got[b].holes) == set(want[b].holes), b
            assert all((got[b][k] == want[b][k]).all() for k in want[b]), b
            assert all((got[b].holes[k] == want[b].holes[k]).all() for k in want[b].holes), b
    no_ring = ssb200.fold_regions_batch({k: v for k, v in tables.items() if k != "ring"}, threads=2)
    assert all(g is None or (not g.holes and all((g[k] == w[k]).all() for k in w)) for g, w in zip(no_ring, want))


@pytest.mark.gpu
def test_frames_to_device_mixed_pinned_and_pageable():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    rng = np.random.default_rng(8)
    shape = (97, 131, 3)
    pinned_src = torch.from_numpy(rng.integers(0, 256, (4,) + shape, dtype=np.uint8)).pin_memory().numpy()
    frames = []
    for i in range(11):
        frames.append(pinned_src[i % 4] if i % 3 == 0 else rng.integers(0, 256, shape, dtype=np.uint8))
    dst = torch.zeros((16,) + shape, dtype=torch.uint8, device="cuda")
    staging = torch.empty((16,) + shape, dtype=torch.uint8).pin_memory()
    stream = torch.cuda.Stream()
    for threads in (1, 6):
        dst.zero_()
        torch.cuda.synchronize()
        ssb200.frames_to_device(frames, dst, staging, threads=threads, stream=stream)
        stream.synchronize()
        assert (dst[:11].cpu().numpy() == np.stack(frames)).all()
        assert int(dst[11:].count_nonzero()) == 0
    # only page-locked frames: no staging buffer needed; pageable ones without it: refused
    ssb200.frames_to_device([pinned_src[1], pinned_src[2]], dst, None, stream=stream)
    stream.synchronize()
    assert (dst[:2].cpu().numpy() == pinned_src[1:3]).all()
    with pytest.raises(ssb200._cabi.SsbError):
        ssb200.frames_to_device([frames[1]], dst, None, stream=stream)
    with pytest.raises(ValueError):
        ssb200.frames_to_device([frames[1][:, :50]], dst, staging, stream=stream)


def test_contour_worker_processes_give_the_in_process_objects():
    """The OpenCV contour stage on worker processes (masks in shared memory) + region tables = the objects of
    the polygon-fill path, in the same order; frames the tables cannot answer for come back as None."""
    import importlib
    import ssb200
    from oracle import segpath as sp
    from test_regions import NAMES, same_objects
    cp = importlib.import_module(ssb200.__name__ + ".contour_pool")
    rng = np.random.default_rng(12)
    H, W, roi_h, roi_w, n = 120, 160, 100, 160, 10
    labs = np.zeros((n, H, W), np.uint8)
    for i in range(n):
        m, l = random_label_map(rng, i)
        hh, ww = min(H, l.shape[0]), min(W, l.shape[1])
        labs[i, :hh, :ww] = l[:hh, :ww]
    labs[0, 10:90, 20:150] = 15
    labs[0, 30:70, 50:120] = 3                      # a hole above the area threshold
    masks = np.zeros_like(labs)
    for i in range(n):
        masks[i, :roi_h, :roi_w] = sp.foreground_mask(labs[i, :roi_h, :roi_w])
    workers = cp.ContourWorkers(3, n * H * W, n_slots=2)
    try:
        workers.slot(1, (n, H, W)).numpy()[:] = masks
        for min_area in (0.0, 40.0):
            workers.submit(1, n, H, W, roi_h, roi_w, min_area)
            cands = workers.collect()
            assert len(cands) == n
            for i in range(n):
                enc = ssb200.fold_region_histograms(*scipy_table(masks[i, :roi_h, :roi_w], labs[i, :roi_h, :roi_w]))
                got = cp.objects_from_candidates(cands[i], enc, NAMES)
                want = ssb200.objects_from_contours(masks[i, :roi_h, :roi_w], labs[i, :roi_h, :roi_w], min_area, NAMES)
                assert got is not None and same_objects(got, want), (i, min_area)
                if cands[i]:
                    assert cp.objects_from_candidates(cands[i], None, NAMES) is None
    finally:
        workers.close()


def _random_masks(rng, n, h, w, rep):
    import cv2
    out = np.zeros((n, h, w), np.uint8)
    for i in range(n):
        kind = (i + rep) % 5
        if kind == 0:
            m = rng.random((h, w)) < rng.uniform(0.1, 0.9)
        elif kind == 1:
            cell = int(rng.integers(2, 7))
            m = np.kron(rng.random((h // cell + 1, w // cell + 1)) < 0.5, np.ones((cell, cell)))[:h, :w] > 0
        elif kind == 2:
            m = np.zeros((h, w), bool)
            for r in range(0, min(h, w) // 2, int(rng.integers(1, 4))):
                m[r:h - r, r:w - r] = r % 2 == 0
            m ^= rng.random((h, w)) < 0.04
        elif kind == 3:
            m = cv2.dilate((rng.random((h, w)) < 0.4).astype(np.uint8), np.ones((2, 2), np.uint8)) > 0
        else:
            m = rng.random((h, w)) >= 0.15
        out[i] = m * 255
    return out


def test_native_contour_stage_equals_opencv_on_ten_thousand_masks():
    """cv2.findContours(RETR_TREE, CHAIN_APPROX_SIMPLE) + contourArea + moments against the library's border
    following: the same contours in the same order, the same outer / hole flags and region keys, and areas and
    centroids equal as doubles / integers -- over noise of every density, blocks, nested rings with specks,
    dilated noise (thin walls) and nearly full frames, at random ROI sizes from 1x1 up."""
    import importlib
    import ssb200
    cw = importlib.import_module(ssb200.__name__ + ".contour_worker")
    rng = np.random.default_rng(5)
    H, W = 48, 64
    frames = contours = 0
    for rep in range(65):
        h, w = int(rng.integers(1, H + 1)), int(rng.integers(1, W + 1))
        masks = np.full((80, H, W), 255, np.uint8)              # whatever lies outside the ROI must not matter
        masks[:, :h, :w] = _random_masks(rng, 80, h, w, rep)
        for min_area in (0.0, 6.5):
            got = ssb200.contour_candidates_batch(masks, h, w, min_area, threads=3, max_candidates=2048)
            for i in range(80):
                want = [c[:5] for c in cw.contour_candidates(masks[i, :h, :w], min_area)]
                assert got[i] == want, (rep, i, h, w, min_area)
                contours += len(want)
            frames += 80
    assert frames >= 10000 and contours > 50000
    # a cut-short candidate list is reported, not silently truncated
    noisy = _random_masks(rng, 2, H, W, 0)
    assert ssb200.contour_candidates_batch(noisy, H, W, 0.0, max_candidates=3) == [None, None]


def _row_moments(mask, first, flags):
    """CPU twin of ssb_contour_moments_u8 for one frame (plain Python): the closed-polygon sums over the border the
    kernel walks for every table row."""
    h, w = mask.shape
    dx, dy = [1, 1, 0, -1, -1, -1, 0, 1], [0, -1, -1, -1, 0, 1, 1, 1]

    def fg(x, y):
        return 0 <= x < w and 0 <= y < h and mask[y, x] != 0
    out = np.zeros((len(first), 3), np.int64)
    for r, (fp, fl) in enumerate(zip(first.tolist(), (flags & 3).tolist())):
        if fl == 2:
            continue
        hole = fl == 0
        y0, x0 = divmod(fp, w)
        x0 -= 1 if hole else 0
        s_end = s = 0 if hole else 4
        while True:
            s = (s - 1) & 7
            if fg(x0 + dx[s], y0 + dy[s]) or s == s_end:
                break
        if not fg(x0 + dx[s], y0 + dy[s]):
            continue
        x1, y1, x, y, pts = x0 + dx[s], y0 + dy[s], x0, y0, []
        while True:
            while True:
                s += 1
                nx, ny = x + dx[s & 7], y + dy[s & 7]
                if fg(nx, ny):
                    break
            s &= 7
            pts.append((x, y))
            if (nx, ny) == (x0, y0) and (x, y) == (x1, y1):
                break
            x, y, s = nx, ny, (s + 4) & 7
        px, py = pts[-1]
        for x, y in pts:
            d = px * y - x * py
            out[r] += (d, d * (px + x), d * (py + y))
            px, py = x, y
    return out


def _row_moments_with_stretches(mask, first, flags):
    """CPU twin of the default kernel behind ssb_contour_moments_u8 (contour_moments_bits_kernel): the same walk, but a
    straight horizontal or vertical stretch is taken in one step -- the walk goes on eastwards from a pixel q iff SW, S,
    SE of q are background and E is foreground (W: NE, N, NW, W; S: NW, W, SW, S; N: SE, E, NE, N), at most 30 pixels at
    a time, never past (x1, y1).  Returns the sums and how many steps were saved."""
    h, w = mask.shape
    dx, dy = [1, 1, 0, -1, -1, -1, 0, 1], [0, -1, -1, -1, 0, 1, 1, 1]

    def fg(x, y):
        return 0 <= x < w and 0 <= y < h and mask[y, x] != 0
    out = np.zeros((len(first), 3), np.int64)
    saved = 0
    for r, (fp, fl) in enumerate(zip(first.tolist(), (flags & 3).tolist())):
        if fl == 2:
            continue
        hole = fl == 0
        y0, x0 = divmod(fp, w)
        x0 -= 1 if hole else 0
        s_end = s = 0 if hole else 4
        while True:
            s = (s - 1) & 7
            if fg(x0 + dx[s], y0 + dy[s]) or s == s_end:
                break
        if not fg(x0 + dx[s], y0 + dy[s]):
            continue
        x1, y1, x, y, pts, travel = x0 + dx[s], y0 + dy[s], x0, y0, [], -1
        while True:
            while True:
                s += 1
                nx, ny = x + dx[s & 7], y + dy[s & 7]
                if fg(nx, ny):
                    break
            s &= 7
            pts.append((x, y))
            if (nx, ny) == (x0, y0) and (x, y) == (x1, y1):
                break
            x, y = nx, ny
            if s == travel and s % 2 == 0:
                ux, uy = dx[s], dy[s]                                  # unit step of the stretch
                # the side that must stay background: below when going east, above (west), left (south), right (north)
                sx, sy = {0: (0, 1), 4: (0, -1), 6: (-1, 0), 2: (1, 0)}[s]
                n = 0
                while n < 30:
                    qx, qy = x + n * ux, y + n * uy
                    side_clear = not any(fg(qx + sx + k * ux, qy + sy + k * uy) for k in (-1, 0, 1))
                    if not (side_clear and fg(qx + ux, qy + uy)):
                        break
                    n += 1
                # never past (x1, y1): the walk may end there
                for k in range(n + 1):
                    if (x + k * ux, y + k * uy) == (x1, y1):
                        n = k
                        break
                x, y = x + n * ux, y + n * uy
                saved += n
            travel = s
            s = (s + 4) & 7
        px, py = pts[-1]
        for x, y in pts:
            d = px * y - x * py
            out[r] += (d, d * (px + x), d * (py + y))
            px, py = x, y
    return out, saved


def test_straight_stretches_in_one_step_leave_the_contour_sums_unchanged():
    """The rule the GPU walk uses to take straight stretches in one step, on noise, blocky masks (long stretches), masks
    whose regions touch the ROI border, one-pixel lines and rings: same a00 / a10 / a01 as the pixel-by-pixel walk."""
    rng = np.random.default_rng(5)
    H, W = 48, 70
    rows = saved = 0
    for rep in range(40):
        if rep % 4 == 0:
            masks = _random_masks(rng, 4, H, W, rep)
        else:
            cell = (2, 3, 5, 8, 13)[rep % 5]
            small = rng.random((4, H // cell + 1, W // cell + 1)) < (0.35 + 0.1 * (rep % 3))
            masks = (np.kron(small, np.ones((1, cell, cell), bool))[:, :H, :W] * 255).astype(np.uint8)
            if rep % 8 == 1:
                masks[:, 0, :] = 255                       # a one-pixel line along the ROI's top row
                masks[:, :, W - 1] = 255
            if rep % 8 == 3:
                masks[0] = 0
                masks[0, 4:44, 5:65] = 255                 # a ring with a one-pixel wall between two holes
                masks[0, 8:40, 9:34] = 0
                masks[0, 8:40, 35:61] = 0
        for i in range(len(masks)):
            n, f, p, fl, _, _ = scipy_table(masks[i], np.zeros_like(masks[i]))
            want = _row_moments(masks[i], f, fl)
            got, s = _row_moments_with_stretches(masks[i], f, fl)
            assert (got == want).all(), (rep, i)
            rows += n
            saved += s
    assert rows > 2000 and saved > 20000, (rows, saved)


def test_candidates_from_tables_and_moments_equal_the_border_following():
    """Region tables + contour integrals -> the candidate list (order, flags, keys, areas, centroids) the border
    following with marks gives on the mask; frames where two holes of one region are candidates ask for the masks."""
    import ssb200
    from oracle import segpath as sp
    rng = np.random.default_rng(31)
    H, W, R = 40, 56, 1024
    checked = asked = 0
    for rep in range(12):
        masks = _random_masks(rng, 16, H, W, rep)
        if rep == 0:                                   # a ring with two big holes separated by a wall, and an island
            masks[0] = 0
            masks[0, 3:37, 4:52] = 255
            masks[0, 8:30, 8:25] = 0
            masks[0, 8:30, 27:48] = 0
            masks[0, 12:20, 30:40] = 255
        labs = np.zeros_like(masks)
        count = np.zeros(16, np.int32)
        first, parent, flags = (np.zeros((16, R), np.int32) for _ in range(3))
        mom = np.zeros((16, R, 3), np.int64)
        for i in range(16):
            n, f, p, fl, _, _ = scipy_table(masks[i], labs[i])
            perm = rng.permutation(n)
            count[i] = n
            first[i, :n], parent[i, :n], flags[i, :n] = f[perm], p[perm], fl[perm]
            mom[i, :n] = _row_moments(masks[i], f[perm], fl[perm])
        tables = {"count": count, "first": first, "parent": parent, "flags": flags, "moments": mom}
        for min_area in (0.5, 30.0, 150.0):
            got = ssb200.candidates_from_moments_batch(tables, W, min_area, threads=3, max_candidates=1024)
            want = ssb200.contour_candidates_batch(masks, H, W, min_area, threads=3, max_candidates=1024)
            for i in range(16):
                if got[i] is None:
                    holes = [c for c in want[i] if c[0]]
                    assert len(holes) >= 2, (rep, i)       # only ever asked for where hole order is at stake
                    asked += 1
                else:
                    assert got[i] == want[i], (rep, i, min_area)
                    checked += 1
    assert checked > 300 and asked >= 1, (checked, asked)


@pytest.mark.gpu
@pytest.mark.parametrize("path", ["bits", "warp"])
def test_gpu_contour_moments_give_the_border_following_candidates(path, monkeypatch):
    """Mask kernel -> region tables -> contour integrals on the GPU -> candidates, against the border following with
    marks on the same masks (which equals OpenCV): every frame that does not ask for the masks must agree exactly.
    Both kernels behind ssb_contour_moments_u8: bit rows in shared memory with a thread per table row (the default),
    and a warp per table row on the mask bytes."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    from oracle import segpath as sp
    monkeypatch.setenv("SSB_MOMENTS_PATH", path)
    rng = np.random.default_rng(77)
    checked = asked = 0
    # 513 x 513: the bit planes need more than the default 48 KB of shared memory; 1100 x 1250: they do not fit at all
    # (the warp-per-row kernel takes over whatever the knob says)
    for (B, H, W, roi_h, roi_w) in ((24, 96, 130, 96, 130), (16, 513, 513, 288, 513), (12, 64, 80, 50, 77), (6, 70, 131, 70, 128),
                                  (5, 40, 35, 33, 32), (3, 513, 513, 513, 513), (2, 1100, 1250, 1100, 1250)):
        labs = np.zeros((B, H, W), np.uint8)
        for i in range(B):
            if i % 3 == 2 or H > 600:
                labs[i] = np.pad(sp.blobby_label_map(i, roi_h, roi_w), ((0, H - roi_h), (0, W - roi_w)))
            else:
                m, l = random_label_map(rng, i)
                reps = (-(-H // l.shape[0]), -(-W // l.shape[1]))
                labs[i] = np.tile(l, reps)[:H, :W]
        lab_d = torch.from_numpy(labs).cuda()
        mask_d = ssb200.foreground_mask(lab_d, roi_h, roi_w)
        t = ssb200.blob_regions(mask_d, lab_d, roi_h, roi_w, 21, 4096)
        ssb200.contour_moments(mask_d, t, roi_h, roi_w)
        torch.cuda.synchronize()
        th = {k: v.cpu().numpy() for k, v in t.items()}
        masks = mask_d.cpu().numpy()
        for min_area in (0.5, 40.0, 0.05 * 513 * 513):
            got = ssb200.candidates_from_moments_batch(th, roi_w, min_area, threads=4, max_candidates=4096)
            want = ssb200.contour_candidates_batch(masks, roi_h, roi_w, min_area, threads=4, max_candidates=4096)
            for i in range(B):
                if got[i] is None:
                    assert th["count"][i] > 4096 or len([c for c in want[i] if c[0]]) >= 2, (B, i)
                    asked += 1
                else:
                    assert got[i] == want[i], (B, H, W, i, min_area)
                    checked += 1
    assert checked > 100


@pytest.mark.gpu
def test_gpu_contour_moments_give_up_on_long_borders(monkeypatch):
    """The walk on the GPU is bounded: a border longer than the step limit marks its row unknown and the frame asks for
    the masks (here with the limit lowered to 40 steps, so that ordinary blobs hit it)."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    from oracle import segpath as sp
    monkeypatch.setenv("SSB_MOMENTS_MAX_STEPS", "40")
    labs = np.stack([np.pad(sp.blobby_label_map(s, 96, 130), ((0, 4), (0, 0))) for s in range(6)]).astype(np.uint8)
    labs[5] = 0
    labs[5, 10:14, 10:14] = 15                           # a border of 12 steps: within the limit
    lab_d = torch.from_numpy(labs).cuda()
    mask_d = ssb200.foreground_mask(lab_d, 96, 130)
    t = ssb200.blob_regions(mask_d, lab_d, 96, 130, 21, 256)
    ssb200.contour_moments(mask_d, t, 96, 130)
    torch.cuda.synchronize()
    th = {k: v.cpu().numpy() for k, v in t.items()}
    got = ssb200.candidates_from_moments_batch(th, 130, 1.0, threads=2)
    want = ssb200.contour_candidates_batch(mask_d.cpu().numpy(), 96, 130, 1.0, threads=2)
    assert got[5] == want[5] and len(want[5]) == 1
    assert sum(g is None for g in got[:5]) >= 3
    assert all(g is None or g == w for g, w in zip(got, want))
