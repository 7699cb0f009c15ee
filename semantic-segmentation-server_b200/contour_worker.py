"""OpenCV contour tail of the producer loop on worker PROCESSES.

`cv2.findContours` does not run in parallel on threads of one process (measured on OpenCV 4.13: 32 masks take
5.2 ms on one thread and 5.0-5.5 ms on two to eight), and at ~80 us per 288x513 mask it alone caps a single
process at ~12 k frames/s before anything else runs.  The reference's contour stage (sem_seg_server.py:88-122:
findContours, contourArea, moments) therefore runs here, in a few plain Python processes that see the masks
through a shared-memory file the parent receives them into straight from the GPU.  A worker returns, per frame
and in OpenCV's contour order, one tuple per contour that passes the area threshold:

    (is_hole, key, area, cx, cy)      key = first pixel (y*w + x) of the region / of the hole the contour
                                      goes around: the row of the GPU's region table that holds what the
                                      contour encloses

This file imports neither torch nor the CUDA library (a worker starts in ~0.3 s) and doubles as the worker's
main program:  python contour_worker.py <shm path> <shm bytes>
"""
import mmap
import os
import pickle
import struct
import sys

import numpy as np


def hole_first_pixel(contour, mask):
    """First pixel (raster order, y*w + x) of the hole a hole contour of `mask` goes around.  The contour
    runs through the foreground pixels 4-adjacent to the hole, so its top row is one above the hole's, and
    a background pixel directly below a contour pixel of that row can only belong to this hole."""
    pts = contour[:, 0, :]
    y_top = int(pts[:, 1].min())
    nxt = np.roll(pts, -1, axis=0)
    on = pts[:, 1] == y_top
    xs = set(int(x) for x in pts[on, 0])
    for a, b in zip(pts[on & (nxt[:, 1] == y_top), 0], nxt[on & (nxt[:, 1] == y_top), 0]):
        xs.update(range(int(min(a, b)), int(max(a, b)) + 1))     # CHAIN_APPROX_SIMPLE keeps only a run's end points
    below = [x for x in sorted(xs) if mask[y_top + 1, x] == 0]
    return (y_top + 1) * mask.shape[1] + below[0] if below else None


def contour_candidates(mask, min_area):
    """sem_seg_server.py:88, :93-94, :115-122 on one mask: contours (RETR_TREE, CHAIN_APPROX_SIMPLE) whose
    polygon area reaches `min_area` and whose m00 is not zero, as (is_hole, key, area, cx, cy) in contour order;
    key is None where the hole of a hole contour could not be identified (the caller fills the polygon then)."""
    import cv2
    mask = np.ascontiguousarray(mask)
    contours, hierarchy = cv2.findContours(mask, cv2.RETR_TREE, cv2.CHAIN_APPROX_SIMPLE)
    out, w = [], mask.shape[1]
    for i in range(len(contours)):
        area = cv2.contourArea(contours[i])
        if area < min_area:
            continue
        m = cv2.moments(contours[i])
        if not m["m00"]:
            continue
        depth, j = 0, hierarchy[0][i][3]
        while j >= 0:                              # outer contours sit at even depth of the RETR_TREE hierarchy
            depth, j = depth + 1, hierarchy[0][j][3]
        if depth % 2 == 0:
            x, y = contours[i][0][0]               # cv2 starts an outer contour at the region's first pixel
            key = int(y) * w + int(x)
        else:
            key = hole_first_pixel(contours[i], mask)
        out.append((depth % 2, key, float(area), int(m["m10"] / m["m00"]), int(m["m01"] / m["m00"]), i))
    return out


def send_msg(f, obj):
    data = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
    f.write(struct.pack("<I", len(data)))
    f.write(data)
    f.flush()


def recv_msg(f):
    head = f.read(4)
    if len(head) < 4:
        return None
    (n,) = struct.unpack("<I", head)
    return pickle.loads(f.read(n))


def worker_main(shm_path, shm_bytes):
    import cv2
    cv2.setNumThreads(1)
    fd = os.open(shm_path, os.O_RDONLY)
    shm = mmap.mmap(fd, shm_bytes, prot=mmap.PROT_READ)
    os.close(fd)
    buf = np.frombuffer(shm, dtype=np.uint8)
    stdin, stdout = sys.stdin.buffer, sys.stdout.buffer
    send_msg(stdout, "ready")
    while True:
        task = recv_msg(stdin)
        if task is None:
            return
        # frames [i0, i1) of a batch of [n][H][W] masks at byte offset `off`, cropped to the ROI
        off, h, w, roi_h, roi_w, i0, i1, min_area = task
        res = []
        for i in range(i0, i1):
            m = buf[off + i * h * w: off + (i + 1) * h * w].reshape(h, w)[:roi_h, :roi_w]
            res.append([c[:5] for c in contour_candidates(m, min_area)])
        send_msg(stdout, res)


if __name__ == "__main__":
    worker_main(sys.argv[1], int(sys.argv[2]))
