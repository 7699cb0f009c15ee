"""Row f1, stage 2: per-contour majority label / purity from the GPU's region table instead of a polygon
fill per contour.  CPU tests check the host logic (tree fold, contour <-> region matching) against the
reference-literal polygon-fill path on tables built with scipy; GPU tests check the kernel's tables
against the same scipy construction and the resulting objects against the reference's fixtures."""
import numpy as np
import pytest

from oracle import segpath as sp  # checker only

ndimage = pytest.importorskip("scipy.ndimage")
S8 = np.ones((3, 3), int)
S4 = np.array([[0, 1, 0], [1, 1, 1], [0, 1, 0]])


def scipy_table(mask, lab, num_classes=21):
    """Region table of one frame as ssb_blob_regions_u8 defines it, rows sorted by first pixel."""
    h, w = mask.shape
    fg = mask > 0
    idx = np.arange(h * w).reshape(h, w)
    rows = []
    region_of = np.empty((h, w), np.int64)          # pixel -> first pixel of its region
    lf, nf = ndimage.label(fg, S8)
    for is_fg, (lbl, n) in ((1, (lf, nf)), (0, ndimage.label(~fg, S4))):
        for k in range(1, n + 1):
            m = lbl == k
            first = int(idx[m].min())
            region_of[m] = first
            touch = (not is_fg) and bool(m[0].any() or m[-1].any() or m[:, 0].any() or m[:, -1].any())
            ring = np.zeros(num_classes, np.int64)
            if not is_fg and not touch:
                y0, x0 = divmod(first, w)            # a hole: the enclosing region owns the pixel left of its first pixel
                ring = np.bincount(lab[ndimage.binary_dilation(m, S4) & (lf == lf[y0, x0 - 1])], minlength=num_classes)
            rows.append((first, is_fg | (2 if touch else 0), np.bincount(lab[m], minlength=num_classes), ring))
    rows.sort(key=lambda r: r[0])
    first = np.array([r[0] for r in rows], np.int32)
    flags = np.array([r[1] for r in rows], np.int32)
    hist = np.array([r[2] for r in rows], np.int32).reshape(len(rows), num_classes)
    ring = np.array([r[3] for r in rows], np.int32).reshape(len(rows), num_classes)
    parent = np.array([-1 if f % w == 0 else region_of.flat[f - 1] for f in first], np.int32)
    return len(rows), first, parent, flags, hist, ring


def random_label_map(rng, t):
    h, w = int(rng.integers(8, 90)), int(rng.integers(8, 120))
    kind = t % 4
    if kind == 0:
        lab = rng.integers(0, 21, (h, w)).astype(np.uint8)
    elif kind == 1:
        cell = int(rng.integers(2, 9))
        small = rng.choice([0, 7, 15, 12], size=(h // cell + 1, w // cell + 1), p=[.4, .25, .25, .1])
        lab = np.kron(small, np.ones((cell, cell), int))[:h, :w].astype(np.uint8)
    elif kind == 2:
        lab = rng.choice([0, 15], size=(h, w), p=[.5, .5]).astype(np.uint8)
    else:
        lab = np.zeros((h, w), np.uint8)
        for r in range(0, min(h, w) // 2, int(rng.integers(2, 5))):
            lab[r:h - r, r:w - r] = 15 if (r // 2) % 2 == 0 else 3
        lab[rng.random((h, w)) < 0.03] = 7
    mask = np.where(lab == 15, 255, 0).astype(np.uint8) if kind == 2 else sp.foreground_mask(lab)
    return mask, lab


def same_objects(a, b):
    return len(a) == len(b) and all(x.label == y.label and x.score == y.score and x.area == y.area and
                                    tuple(x.centroid) == tuple(y.centroid) for x, y in zip(a, b))


NAMES = {i: "label%d" % i for i in range(21)}


def test_objects_from_region_tables_equal_the_polygon_fill_path():
    """Noise, blocks with holes, pixel-level diagonal patterns and nested rings; no area threshold, so
    every contour (also hole contours, which take the fill path inside) is compared."""
    import ssb200
    rng = np.random.default_rng(4)
    n_outer = n_holes = 0

    def no_label_map():
        raise AssertionError("the tables cover every contour: the label map must not be needed")
    for t in range(120):
        mask, lab = random_label_map(rng, t)
        count, first, parent, flags, hist, ring = scipy_table(mask, lab)
        enclosed = ssb200.fold_region_histograms(count, first, parent, flags, hist, ring)
        n_outer += len(enclosed)
        n_holes += len(enclosed.holes)
        want = ssb200.objects_from_contours(mask, lab, 0.0, NAMES)
        assert same_objects(ssb200.objects_from_regions(mask, enclosed, 0.0, NAMES, label_map=no_label_map), want), t
        # without ring rows hole contours take the polygon fill, with the same result
        no_ring = ssb200.fold_region_histograms(count, first, parent, flags, hist)
        assert same_objects(ssb200.objects_from_regions(mask, no_ring, 0.0, NAMES, label_map=lab), want), t
        # rows in any order (the kernel hands out rows with an atomic counter)
        perm = rng.permutation(count)
        enclosed2 = ssb200.fold_region_histograms(count, first[perm], parent[perm], flags[perm], hist[perm], ring[perm])
        assert all((enclosed[k] == enclosed2[k]).all() for k in enclosed)
        assert all((enclosed.holes[k] == enclosed2.holes[k]).all() for k in enclosed.holes)
    assert n_outer > 500 and n_holes > 2000


def test_incomplete_table_falls_back_to_the_fill(golden):
    import ssb200
    g = golden("objects.npz")
    name = sorted(k[:-5] for k in g.files if k.endswith("_geom"))[0]
    cw, ch, W, H = (int(v) for v in g[name + "_geom"][:4])
    geom = ssb200.FrameGeometry(cw, ch, W, H)
    lab = np.ascontiguousarray(g[name + "_label_map"][:geom.roi_h, :geom.roi_w]).astype(np.uint8)
    mask = sp.foreground_mask(lab)
    count, first, parent, flags, hist, ring = scipy_table(mask, lab)
    assert count >= 2
    assert ssb200.fold_region_histograms(count, first[:1], parent[:1], flags[:1], hist[:1]) is None
    fetched = []
    got = ssb200.objects_from_regions(mask, None, 100.0, NAMES, label_map=lambda: fetched.append(1) or lab)
    assert same_objects(got, ssb200.objects_from_contours(mask, lab, 100.0, NAMES)) and len(got) >= 1
    assert len(fetched) == 1                          # the label map is fetched once, and only because it was needed
    enclosed = ssb200.fold_region_histograms(count, first, parent, flags, hist)
    got = ssb200.objects_from_regions(mask, enclosed, 100.0, NAMES, label_map=lambda: fetched.append(1) or lab)
    assert same_objects(got, ssb200.objects_from_contours(mask, lab, 100.0, NAMES))


# ---- GPU ------------------------------------------------------------------------------------------------
def gpu_tables(ssb, masks, labs, roi_h, roi_w, max_regions):
    import torch
    t = ssb.blob_regions(torch.from_numpy(masks).cuda(), torch.from_numpy(labs).cuda(), roi_h, roi_w, 21, max_regions)
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in t.items()}


@pytest.mark.gpu
@pytest.mark.parametrize("path", ["cluster", "global"])
def test_gpu_region_tables_match_scipy_labelling(path, monkeypatch):
    """Both implementations behind ssb_blob_regions_u8: one thread-block cluster per frame with the labelling in
    shared memory (the default for ROIs of this size), and the global-memory kernels that take any ROI."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    monkeypatch.setenv("SSB_BLOB_PATH", path)
    rng = np.random.default_rng(9)
    H, W = 96, 130
    for roi_h, roi_w in ((96, 130), (57, 130), (96, 77), (40, 33)):
        masks = np.zeros((8, H, W), np.uint8)
        labs = np.zeros((8, H, W), np.uint8)
        for b in range(8):
            m, l = random_label_map(rng, b)
            hh, ww = min(m.shape[0], H), min(m.shape[1], W)
            masks[b, :hh, :ww], labs[b, :hh, :ww] = m[:hh, :ww], l[:hh, :ww]
        labs[:, :, :] = np.where(labs < 21, labs, 0)
        t = gpu_tables(ssb200, masks, labs, roi_h, roi_w, 8192)
        for b in range(8):
            count, first, parent, flags, hist, ring = scipy_table(masks[b, :roi_h, :roi_w], labs[b, :roi_h, :roi_w])
            assert t["count"][b] == count
            order = np.argsort(t["first"][b, :count])
            assert (t["first"][b, :count][order] == first).all()
            assert (t["parent"][b, :count][order] == parent).all()
            assert (t["flags"][b, :count][order] == flags).all()
            assert (t["hist"][b, :count][order] == hist).all()
            holes = (flags & 3) == 0                     # ring rows are defined for holes only
            assert (t["ring"][b, :count][order][holes] == ring[holes]).all()


@pytest.mark.gpu
def test_gpu_regions_reproduce_reference_objects(golden):
    """The reference's own scenes: mask kernel -> region kernel -> objects, against what the reference's
    loop left in its deque (same fixtures as the polygon-fill path)."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    g = golden("objects.npz")
    lab_names = golden("labels_colormap.npz")
    names = dict(zip(lab_names["label_ids"].tolist(), lab_names["label_names"].tolist()))
    checked = 0
    for name in sorted(k[:-5] for k in g.files if k.endswith("_geom")):
        cw, ch, W, H = (int(v) for v in g[name + "_geom"][:4])
        geom = ssb200.FrameGeometry(cw, ch, W, H)
        lm = torch.from_numpy(np.ascontiguousarray(g[name + "_label_map"])[None].astype(np.uint8)).cuda()
        mask = ssb200.foreground_mask(lm, geom.roi_h, geom.roi_w)
        t = ssb200.blob_regions(mask, lm, geom.roi_h, geom.roi_w, 21, 256)
        torch.cuda.synchronize()
        th = {k: v.cpu().numpy()[0] for k, v in t.items()}
        enclosed = ssb200.fold_region_histograms(th["count"], th["first"], th["parent"], th["flags"], th["hist"], th["ring"])
        mask_h = mask[0, :geom.roi_h, :geom.roi_w].cpu().numpy()
        lab_h = lm[0, :geom.roi_h, :geom.roi_w].cpu().numpy()
        objs = ssb200.objects_from_regions(mask_h, enclosed, float(g[name + "_min_area_ratio"]) * W * H, names,
                                           label_map=lambda: lab_h)
        vals = np.array([list(ssb200.normalise(o, W, H)[1:]) for o in objs], dtype=np.float32).reshape(-1, 4)
        assert [o.label for o in objs][::-1] == g[name + "_obj_names"].tolist(), name      # fixtures are newest-first
        assert (vals[::-1] == g[name + "_obj_vals"]).all(), name
        checked += len(objs)
    assert checked >= 8


@pytest.mark.gpu
def test_gpu_large_hole_contours_need_no_label_map():
    """A ring with a hole above the area threshold, an island in the hole and a hole in the island: outer and
    hole contours all come from the tables (the reference reports the big hole as an object of its own)."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    lab = np.zeros((2, 513, 513), np.uint8)
    lab[0, 20:270, 30:480] = 15
    lab[0, 60:230, 90:420] = 3              # hole: 170 x 330 = 56 100 px
    lab[0, 100:190, 150:330] = 7            # island in the hole (90 x 180 = 16 200 px)
    lab[0, 120:160, 180:300] = 0            # hole in the island
    lab[0, 62, 95:100] = 15                 # specks in the big hole
    lab[1, 5:280, 5:508] = 7
    lab[1, 40:240, 40:200] = 12
    lab[1, 40:240, 201:470] = 9             # two holes separated by a one-pixel wall
    masks = np.zeros_like(lab)
    for b in range(2):
        masks[b, :288] = sp.foreground_mask(lab[b, :288])
    t = gpu_tables(ssb200, masks, lab, 288, 513, 256)

    def no_label_map():
        raise AssertionError("label map requested")
    n_obj = 0
    for b in range(2):
        enclosed = ssb200.fold_region_histograms(t["count"][b], t["first"][b], t["parent"][b], t["flags"][b], t["hist"][b],
                                                 t["ring"][b])
        got = ssb200.objects_from_regions(masks[b, :288], enclosed, 0.05 * 513 * 513, NAMES, label_map=no_label_map)
        want = ssb200.objects_from_contours(masks[b, :288], lab[b, :288], 0.05 * 513 * 513, NAMES)
        assert same_objects(got, want), b
        n_obj += len(got)
    assert n_obj >= 5


@pytest.mark.gpu
def test_gpu_regions_full_size_noise_and_overflow():
    """288x513 ROI of a 513x513 frame: blobby maps fit the table, pure noise overflows it (count says so)."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    labs = np.stack([np.pad(sp.blobby_label_map(s, 288, 513), ((0, 225), (0, 0))) for s in range(3)] +
                    [np.random.default_rng(1).integers(0, 21, (513, 513)).astype(np.uint8)]).astype(np.uint8)
    masks = np.zeros_like(labs)
    for b in range(4):
        masks[b, :288] = sp.foreground_mask(labs[b, :288])
    t = gpu_tables(ssb200, masks, labs, 288, 513, 256)
    for b in range(3):
        enclosed = ssb200.fold_region_histograms(t["count"][b], t["first"][b], t["parent"][b], t["flags"][b], t["hist"][b],
                                                 t["ring"][b])
        got = ssb200.objects_from_regions(masks[b, :288], enclosed, 0.05 * 513 * 513, NAMES, label_map=labs[b, :288])
        assert same_objects(got, ssb200.objects_from_contours(masks[b, :288], labs[b, :288], 0.05 * 513 * 513, NAMES))
        assert int(t["hist"][b, :t["count"][b]].sum()) == 288 * 513
    count, *_ = scipy_table(masks[3, :288], labs[3, :288])
    assert t["count"][3] == count
    if count > 256:
        assert ssb200.fold_region_histograms(t["count"][3], t["first"][3], t["parent"][3], t["flags"][3], t["hist"][3]) is None


@pytest.mark.gpu
def test_gpu_regions_wide_roi_and_many_rows():
    """Shapes around the limits of the cluster kernel: an ROI whose strips are wider than one warp-word row
    count (1 x N, N x 1, 3 x 700), more table rows than fit 16 bits (-> global-memory kernels), a ROI narrower
    than the frame (no staged copy), and a batch of frames with different contents."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    rng = np.random.default_rng(21)
    for (H, W, roi_h, roi_w, rows) in ((5, 700, 3, 700, 512), (300, 9, 300, 1, 64), (40, 64, 1, 64, 64),
                                       (130, 200, 129, 200, 70000), (200, 300, 200, 300, 4096), (64, 96, 64, 95, 300)):
        labs = rng.choice([0, 7, 15, 3], size=(5, H, W), p=[.45, .2, .25, .1]).astype(np.uint8)
        labs[0] = 15
        labs[1] = 0
        masks = np.zeros_like(labs)
        for b in range(5):
            masks[b, :roi_h, :roi_w] = sp.foreground_mask(labs[b, :roi_h, :roi_w])
        t = gpu_tables(ssb200, masks, labs, roi_h, roi_w, rows)
        for b in range(5):
            count, first, parent, flags, hist, ring = scipy_table(masks[b, :roi_h, :roi_w], labs[b, :roi_h, :roi_w])
            assert t["count"][b] == count, (H, W, b)
            if count > rows:
                continue
            order = np.argsort(t["first"][b, :count])
            for name, want in (("first", first), ("parent", parent), ("flags", flags), ("hist", hist)):
                assert (t[name][b, :count][order] == want).all(), (H, W, b, name)
            holes = (flags & 3) == 0
            assert (t["ring"][b, :count][order][holes] == ring[holes]).all(), (H, W, b)


@pytest.mark.gpu
@pytest.mark.parametrize("path", ["cluster", "global"])
def test_gpu_regions_uniform_stretches_at_every_byte_phase(path, monkeypatch):
    """Label maps made of large one-label cells (words that are counted whole by the cluster kernel), with single odd
    pixels inside the stretches, rows whose last word is short, and label maps that start at byte phases 0..3."""
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ssb200
    monkeypatch.setenv("SSB_BLOB_PATH", path)
    rng = np.random.default_rng(33)
    for case, (H, W, roi_h, roi_w, cell) in enumerate(((150, 257, 150, 257, 40), (96, 330, 90, 301, 64), (130, 160, 130, 160, 16),
                                                       (77, 515, 77, 515, 100), (64, 96, 64, 96, 32), (120, 131, 120, 131, 24),
                                                       (728, 720, 728, 720, 56))):     # more words per strip than the shared list holds
        B = 4
        cells = rng.choice([0, 15, 7, 3, 12], size=(B, H // cell + 2, W // cell + 2), p=[.3, .35, .15, .1, .1]).astype(np.uint8)
        labs = np.repeat(np.repeat(cells, cell, 1), cell, 2)[:, 3:H + 3, 5:W + 5].copy()
        labs[:, ::17, ::23] = 7
        labs[:, 5::29, 7::31] = 0
        labs[3] = 15 if case % 2 else 0
        masks = np.zeros_like(labs)
        for b in range(B):
            masks[b, :roi_h, :roi_w] = sp.foreground_mask(labs[b, :roi_h, :roi_w])
        phase = case % 4
        buf = torch.zeros(B * H * W + 8, dtype=torch.uint8, device="cuda")
        lab_d = buf[phase:phase + B * H * W].view(B, H, W)
        lab_d.copy_(torch.from_numpy(labs))
        assert lab_d.data_ptr() % 4 == phase
        t = ssb200.blob_regions(torch.from_numpy(masks).cuda(), lab_d, roi_h, roi_w, 21, 2048)
        torch.cuda.synchronize()
        t = {k: v.cpu().numpy() for k, v in t.items()}
        for b in range(B):
            count, first, parent, flags, hist, ring = scipy_table(masks[b, :roi_h, :roi_w], labs[b, :roi_h, :roi_w])
            assert t["count"][b] == count, (case, b)
            order = np.argsort(t["first"][b, :count])
            for name, want in (("first", first), ("parent", parent), ("flags", flags), ("hist", hist)):
                assert (t[name][b, :count][order] == want).all(), (case, b, name)
            holes = (flags & 3) == 0
            assert (t["ring"][b, :count][order][holes] == ring[holes]).all(), (case, b)
            assert int(t["hist"][b, :count].sum()) == roi_h * roi_w
