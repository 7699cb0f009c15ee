"""Drop-in for the reference's server module (sem_seg_server.py): same gRPC service, same flags, same
reply semantics, with the per-frame path running on the GPU in batches.

    recognize_and_segment   sem_seg_server.py:135-213   producer loop (batched)
    SemanticSegmentationServicer  :216-234              LIFO pop x num_detections, padded with empties
    serve                   :236-291                    flags, engine/labels/camera set-up, gRPC lifecycle

The model (`engine`) is a duck type, as in the reference: `get_input_tensor_shape()` -> (1,H,W,3) and
`run_logits(input_u8 [B,H,W,3] cuda) -> (logits [B,H,W,C] cuda u8|f16, scale, zero_point)`, or planar
[B,C,H,W] f32|f16 scores when the engine sets `layout = "nchw"`.  The Edge TPU tflite model is
unavailable here; `SyntheticEngine` produces class scores of the model's exact output shape and
quantisation, `TorchvisionEngine` runs a DeepLabv3 network on the GPU.
"""
import argparse
import collections
import concurrent.futures
import os

import numpy as np
import torch

from . import contour_pool, hotpath, proto

# Stack (LIFO deque) shared by the producer thread and the RPC handlers (sem_seg_server.py:28).
segmented_objects = collections.deque()


class SyntheticEngine:
    """Stands in for BasicEngine (sem_seg_server.py:264): quantised class scores derived from the
    input image so that results are deterministic and depend on the frame."""

    def __init__(self, height=513, width=513, num_classes=21, scale=1.0 / 16, zero_point=128):
        self.h, self.w, self.c, self.scale, self.zero_point = height, width, num_classes, scale, zero_point

    def get_input_tensor_shape(self):
        return (1, self.h, self.w, 3)

    def run_logits(self, x):
        b = x.shape[0]
        lum = x.to(torch.int32).sum(dim=-1)                                    # [B,H,W]
        ch = torch.arange(self.c, device=x.device, dtype=torch.int32)
        q = ((lum.unsqueeze(-1) * (ch * 7 + 3) + ch * 31) % 256).to(torch.uint8)
        return q.contiguous().view(b, self.h, self.w, self.c), self.scale, self.zero_point


class TorchvisionEngine:
    """A logits provider with a real network behind it: torchvision's
    DeepLabv3-MobileNetV3 with 21 classes, randomly initialised (no checkpoint can be fetched here; pass
    `state_dict` to load one).  It hands over what the head leaves on the GPU -- planar [B,C,H,W] fp32
    scores, fp16 with `half=True` -- so the argmax + statistics pass reads it in place
    (`layout = "nchw"` selects hotpath.argmax_stats_nchw in recognize_and_segment)."""

    layout = "nchw"

    def __init__(self, height=513, width=513, num_classes=21, half=False, device="cuda", seed=0, state_dict=None):
        import torchvision
        self.h, self.w, self.c, self.half = height, width, num_classes, half
        torch.manual_seed(seed)
        self.model = torchvision.models.segmentation.deeplabv3_mobilenet_v3_large(
            weights=None, weights_backbone=None, num_classes=num_classes).eval().to(device)
        if state_dict is not None:
            self.model.load_state_dict(state_dict)
        self.mean = torch.tensor([0.485, 0.456, 0.406], device=device).view(1, 3, 1, 1) * 255.0
        self.std = torch.tensor([0.229, 0.224, 0.225], device=device).view(1, 3, 1, 1) * 255.0

    def get_input_tensor_shape(self):
        return (1, self.h, self.w, 3)

    @torch.no_grad()
    def run_logits(self, x):
        """x: [B,H,W,3] u8 RGB (the output of the pre-process kernel) -> ([B,C,H,W] scores, 1.0, 0)."""
        inp = (x.permute(0, 3, 1, 2).float() - self.mean) / self.std
        if self.half:
            with torch.autocast(device_type=inp.device.type, dtype=torch.float16):
                out = self.model(inp)["out"]
            out = out.to(torch.float16)
        else:
            out = self.model(inp)["out"]
        return out.contiguous(), 1.0, 0


class ArrayCamera:
    """cv2.VideoCapture look-alike over an in-memory list of BGR frames (tests, benchmarks)."""

    def __init__(self, frames):
        self._frames, self._i = list(frames), 0

    def isOpened(self):
        return True

    def read(self):
        if self._i >= len(self._frames):
            return False, None
        self._i += 1
        return True, self._frames[self._i - 1]

    def release(self):
        pass

    def get(self, prop):
        import cv2
        h, w = self._frames[0].shape[:2]
        return {cv2.CAP_PROP_FRAME_WIDTH: w, cv2.CAP_PROP_FRAME_HEIGHT: h}.get(prop, 0)


REGION_ROWS = 256     # rows of the per-frame region table; frames with more regions use the polygon fill
HOST_THREADS = min(16, len(os.sched_getaffinity(0)))   # host threads of the native helpers (frame gather, table fold)

CONTOUR_WORKERS = int(os.environ.get("SSB_CONTOUR_WORKERS", min(8, max(1, len(os.sched_getaffinity(0)) // 2))))
_pool = None
_workers = None


def _contour_workers(n_workers, slot_bytes):
    """The process-wide contour worker pool (started on first use, restarted when a larger batch arrives)."""
    global _workers
    if not n_workers or n_workers <= 0:
        return None
    if _workers is None or _workers.slot_bytes < slot_bytes or len(_workers._procs) != n_workers:
        if _workers is not None:
            _workers.close()
        _workers = contour_pool.ContourWorkers(n_workers, slot_bytes)
        import atexit
        atexit.register(_workers.close)
    return _workers


def _contour_pool():
    global _pool
    if _pool is None:
        _pool = concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(os.sched_getaffinity(0))))
    return _pool


_buffers = {}


def _feeder_buffers(dev, slot, shape):
    """Page-locked staging + device input buffer of a feeder slot, kept for the life of the process (locking
    hundreds of MB takes ~0.1 s: not something to repeat per call of recognize_and_segment)."""
    key = (str(dev), slot, tuple(shape))
    if key not in _buffers:
        for k in [k for k in _buffers if k[:2] == key[:2]]:      # another frame size took this slot before
            del _buffers[k]
        _buffers[key] = (torch.empty(shape, dtype=torch.uint8).pin_memory(), torch.empty(shape, dtype=torch.uint8, device=dev))
    return _buffers[key]


def _open_camera(camera_idx):
    if hasattr(camera_idx, "read"):
        return camera_idx
    import cv2
    return cv2.VideoCapture(camera_idx)


def _feed_batches(cap, engine, keep_aspect_ratio, width, height, batch_frames, mode, dev, out_q, stop, gpu_contours=False):
    """Feeder thread of recognize_and_segment: reads frames, gathers them into one of two pinned buffers,
    and enqueues transfer + K1 + model + K2/K3 (+ K4/K5) on its own CUDA stream.  Each batch goes to the
    consumer as (n_frames, geometry, device results, CUDA event); None ends the stream, an exception is
    passed through."""
    try:
        stream = torch.cuda.Stream(dev)
        pinned, busy = [None, None], [None, None]      # two frame buffers and the event of their last transfer
        device_in, held, consumed = [None, None], [None, None], [None, None]
        up_stream = torch.cuda.Stream(dev)
        geom, k = None, 0
        tables = [None] * 4
        while cap.isOpened() and not stop.is_set():
            frames = []
            while len(frames) < batch_frames:
                ret, frame = cap.read()
                if not ret:
                    break
                frames.append(frame)
            if not frames:
                break
            cam_h, cam_w = frames[0].shape[:2]
            if geom is None or (geom.cam_w, geom.cam_h) != (cam_w, cam_h):
                geom = hotpath.FrameGeometry(cam_w, cam_h, width, height, keep_aspect_ratio)
            slot = k % 2
            if busy[slot] is not None:
                busy[slot].synchronize()               # the transfer that last read this slot's buffers has finished
            if pinned[slot] is None or pinned[slot].shape[1:] != (cam_h, cam_w, 3):
                pinned[slot], device_in[slot] = _feeder_buffers(dev, slot, (batch_frames, cam_h, cam_w, 3))
            frames = [f if (f.dtype == np.uint8 and f.flags.c_contiguous) else np.ascontiguousarray(f, np.uint8) for f in frames]
            held[slot] = frames                        # frames in page-locked memory are read by the DMA where they are
            # Uploads have a stream of their own: on the kernels' stream the copy of batch k+1 would wait for the
            # kernels of batch k.  It waits only for the pre-process kernel that last read this slot's device buffer.
            if consumed[slot] is not None:
                up_stream.wait_event(consumed[slot])
            # native: pageable frames are gathered into the pinned buffer by host threads and sent frame by
            # frame (gather and DMA overlap), page-locked ones go straight from their own memory
            batch = hotpath.frames_to_device(frames, device_in[slot], pinned[slot], threads=HOST_THREADS,
                                             stream=up_stream)[:len(frames)]
            busy[slot] = torch.cuda.Event()
            busy[slot].record(up_stream)
            with torch.cuda.stream(stream):
                stream.wait_event(busy[slot])
                inputs = hotpath.preprocess_frames(batch, geom)                   # :150-161
                consumed[slot] = torch.cuda.Event()
                consumed[slot].record(stream)
                logits, scale, zero_point = engine.run_logits(inputs)             # :162 (model, upstream)
                nchw = getattr(engine, "layout", "nhwc") == "nchw"
                if nchw:                                                          # :163-170, :102-122
                    label_maps, stats = hotpath.argmax_stats_nchw(logits, geom.roi_h, geom.roi_w)
                else:
                    label_maps, stats = hotpath.argmax_stats(logits, geom.roi_h, geom.roi_w)
                res = {"label_maps": label_maps, "stats": stats, "scale": scale, "zero_point": zero_point,
                       "float_scores": logits.dtype != torch.uint8}
                if mode == "contours":
                    masks = hotpath.foreground_mask(label_maps, geom.roi_h, geom.roi_w)   # :173, :79-85
                    # :98-111 for every contour at once: region tables instead of one polygon fill per contour
                    # the table tensors rotate over four sets: one being filled, one queued, one being read back
                    tables[k % 4] = hotpath.blob_regions(masks, label_maps, geom.roi_h, geom.roi_w,
                                                         logits.shape[1 if nchw else 3], REGION_ROWS, out=tables[k % 4])
                    if gpu_contours:
                        hotpath.contour_moments(masks, tables[k % 4], geom.roi_h, geom.roi_w)   # :93, :115 as exact integers
                    res["tables"] = dict(tables[k % 4], mask=masks)
                done = torch.cuda.Event()
                done.record(stream)
            out_q.put((len(frames), geom, res, done))
            k += 1
        out_q.put(None)
    except BaseException as exc:   # handed to the consumer, which re-raises it
        out_q.put(exc)


def recognize_and_segment(keep_aspect_ratio, camera_idx, engine, labels, min_area_ratio,
                          batch_frames=16, mode="contours", device=0, max_frames=None, contour_workers=None):
    """Recognize and segment objects from camera frames (sem_seg_server.py:135-213), `batch_frames`
    frames per GPU pass.  `mode`: "contours" reproduces the reference's objects (per contour of the
    thresholded colour image); "labels" emits the per-label objects of the fused statistics.

    Two stages run side by side: a feeder thread gathers batch n+1 and enqueues its GPU work while this
    thread reads back batch n, runs the OpenCV tails and pushes the objects -- in frame order, so the deque
    looks exactly as the reference's.  `contour_workers`: where the contour stage (sem_seg_server.py:88-122)
    runs -- None or "gpu": contour integrals on the GPU with the region tables (masks stay on the device; the rare
    frame whose contour order cannot be told from the tables, or whose border exceeds the walk limit, goes through the
    host border following); "host": the library's own border following on host threads for every frame (same contours
    and numbers as OpenCV); k > 0: cv2 in k worker processes; 0: cv2 in this process."""
    import queue
    import threading
    if contour_workers is None:
        contour_workers = "gpu"
    segmented_objects.clear()
    _, height, width, _ = engine.get_input_tensor_shape()
    segment_centroid_min_area = min_area_ratio * height * width
    dev = torch.device("cuda", device)
    cap = _open_camera(camera_idx)
    done = n_batches = 0
    staging = hotpath.HostStaging()
    fold_scratch = {}
    batches, stop = queue.Queue(maxsize=1), threading.Event()
    feeder = threading.Thread(target=_feed_batches, daemon=True,
                              args=(cap, engine, keep_aspect_ratio, width, height, batch_frames, mode, dev, batches, stop,
                                    contour_workers == "gpu"))
    feeder.start()
    copy_stream = torch.cuda.Stream(dev)
    try:
        while True:
            item = batches.get()
            if item is None:
                break
            if isinstance(item, BaseException):
                raise item
            n_frames, geom, res, ready = item
            workers = (_contour_workers(contour_workers, batch_frames * height * width)
                       if mode == "contours" and isinstance(contour_workers, int) and contour_workers > 0 else None)
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ready)
                if mode == "contours" and workers is not None:
                    # masks go by DMA into the shared-memory slot the worker processes read; tables to pinned buffers
                    slot = n_batches % workers.n_slots
                    masks_sh = workers.slot(slot, (n_frames, height, width))
                    masks_sh.copy_(res["tables"]["mask"][:n_frames], non_blocking=True)
                    tables_h = staging.fetch({k: v for k, v in res["tables"].items() if k != "mask"})   # one wait for all
                    tables_h["mask"] = masks_sh.numpy()
                elif mode == "contours" and contour_workers == "gpu":
                    # the contour integrals came from the GPU with the tables: the masks stay on the device
                    tables_h = staging.fetch({k: v for k, v in res["tables"].items() if k != "mask"})
                elif mode == "contours":
                    tables_h = staging.fetch(res["tables"])                       # pinned buffers, one wait
                else:
                    stats_h = staging.fetch({"stats": res["stats"]})["stats"]
            n_batches += 1
            if mode == "contours" and contour_workers == "gpu":
                enclosed = hotpath.fold_regions_batch(tables_h, HOST_THREADS, fold_scratch)
                candidates = hotpath.candidates_from_moments_batch(tables_h, geom.roi_w, segment_centroid_min_area, HOST_THREADS,
                                                                   scratch=fold_scratch)
                per_frame = []
                for i in range(n_frames):
                    objs = None if candidates[i] is None else contour_pool.objects_from_candidates(candidates[i], enclosed[i], labels)
                    if objs is None:
                        # incomplete table, or the order of two candidate holes of one region is at stake: this frame's
                        # mask comes to the host after all and goes through the border following with marks
                        mask_i = res["tables"]["mask"][i:i + 1].cpu().numpy()
                        cand = hotpath.contour_candidates_batch(mask_i, geom.roi_h, geom.roi_w, segment_centroid_min_area, 1)[0]
                        objs = None if cand is None else contour_pool.objects_from_candidates(cand, enclosed[i], labels)
                    if objs is None:
                        objs = hotpath.objects_from_regions(
                            res["tables"]["mask"][i, :geom.roi_h, :geom.roi_w].cpu().numpy(), enclosed[i], segment_centroid_min_area,
                            labels, label_map=lambda: res["label_maps"][i, :geom.roi_h, :geom.roi_w].cpu().numpy())
                    per_frame.append(objs)
            elif mode == "contours":
                masks_h = tables_h["mask"][:, :geom.roi_h, :geom.roi_w]
                label_maps = res["label_maps"]
                if workers is not None:
                    workers.submit(slot, n_frames, height, width, geom.roi_h, geom.roi_w, segment_centroid_min_area)
                # what every contour encloses, for the whole batch at once (native, HOST_THREADS threads)
                enclosed = hotpath.fold_regions_batch(tables_h, HOST_THREADS, fold_scratch)

                def contour_tail(i):
                    # the label map is only needed for frames with more regions than table rows: fetched on demand
                    return hotpath.objects_from_regions(
                        masks_h[i], enclosed[i], segment_centroid_min_area, labels,
                        label_map=lambda: label_maps[i, :geom.roi_h, :geom.roi_w].cpu().numpy())
                if contour_workers == 0:
                    per_frame = list(_contour_pool().map(contour_tail, range(n_frames)))     # cv2, in this process
                else:
                    if workers is not None:
                        # cv2.findContours / contourArea / moments ran in the worker processes
                        candidates = workers.collect()
                    else:
                        # the same contours, areas and centroids from the library's border following (equal to
                        # OpenCV's, tests/test_hostpath.py), a frame per host thread
                        candidates = hotpath.contour_candidates_batch(tables_h["mask"][:n_frames], geom.roi_h, geom.roi_w,
                                                                      segment_centroid_min_area, HOST_THREADS, scratch=fold_scratch)
                    # the candidates meet the region tables, in frame order
                    per_frame = [None if candidates[i] is None else
                                 contour_pool.objects_from_candidates(candidates[i], enclosed[i], labels) for i in range(n_frames)]
                    per_frame = [contour_tail(i) if objs is None else objs for i, objs in enumerate(per_frame)]
            else:
                shift = hotpath._cabi.F16_FIXED_SHIFT if res["float_scores"] else 0
                per_frame = [hotpath.objects_from_stats(stats_h[i], width, height, labels, min_area_ratio, res["scale"],
                                                        res["zero_point"], fixed_shift=shift)[0] for i in range(n_frames)]
            for objs in per_frame:
                for o in objs:
                    lab, score, area, cx, cy = hotpath.normalise(o, width, height)  # :184-192
                    segmented_objects.appendleft(proto.SegmentedObject(
                        label=lab, score=score, area=area, centroid=proto.SegmentedObject.Centroid(cx=cx, cy=cy)))
            done += n_frames
            if max_frames is not None and done >= max_frames:
                break
    except KeyboardInterrupt:
        pass
    finally:
        stop.set()
        while feeder.is_alive():                       # unblock a feeder waiting to hand over a batch
            try:
                batches.get(timeout=0.05)
            except queue.Empty:
                pass
        feeder.join()
        cap.release()
    return


class SemanticSegmentationServicer:
    """sem_seg_server.py:216-234, registered through generic handlers (no generated grpc module)."""

    def __init__(self, num_detections, camera_res):
        self.num_detections = num_detections
        self.camera_res = camera_res

    def GetCameraResolution(self, request, context):
        return proto.CameraResolution(width=self.camera_res[0], height=self.camera_res[1])

    def GetSegmentedObjects(self, request, context):
        def pop_deque():
            try:
                return segmented_objects.popleft()
            except IndexError:
                return proto.SegmentedObject()      # stack empty: empty object (:229-231)
        return proto.SegmentedObjectData(data=[pop_deque() for _ in range(self.num_detections)])


def add_SemanticSegmentationServicer_to_server(servicer, server):
    """Same registration as sem_seg_server_pb2_grpc.py:48-63."""
    import grpc
    handlers = {
        "GetSegmentedObjects": grpc.unary_unary_rpc_method_handler(
            servicer.GetSegmentedObjects, request_deserializer=proto.Empty.FromString,
            response_serializer=proto.SegmentedObjectData.SerializeToString),
        "GetCameraResolution": grpc.unary_unary_rpc_method_handler(
            servicer.GetCameraResolution, request_deserializer=proto.Empty.FromString,
            response_serializer=proto.CameraResolution.SerializeToString),
    }
    server.add_generic_rpc_handlers(
        (grpc.method_handlers_generic_handler("%s.%s" % (proto.PACKAGE, proto.SERVICE), handlers),))


class SemanticSegmentationStub:
    """Client stub equivalent to sem_seg_server_pb2_grpc.py:7-26."""

    def __init__(self, channel):
        self.GetSegmentedObjects = channel.unary_unary(
            proto.METHOD_GET_OBJECTS, request_serializer=proto.Empty.SerializeToString,
            response_deserializer=proto.SegmentedObjectData.FromString)
        self.GetCameraResolution = channel.unary_unary(
            proto.METHOD_GET_RESOLUTION, request_serializer=proto.Empty.SerializeToString,
            response_deserializer=proto.CameraResolution.FromString)


def build_arg_parser():
    default_model_dir = "/media/mendel/sem-seg-server/models"
    parser = argparse.ArgumentParser()
    parser.add_argument("--model", help=".tflite model path (unused: synthetic class scores)",
                        default=os.path.join(default_model_dir, "deeplabv3_mnv2_pascal_quant_edgetpu.tflite"))
    parser.add_argument("--labels", help="label file path",
                        default=os.path.join(default_model_dir, "pascal_voc_segmentation_labels.txt"))
    parser.add_argument("--keep_aspect_ratio", dest="keep_aspect_ratio", action="store_true",
                        help="keep the image aspect ratio when down-sampling by adding zero padding on bottom or right")
    parser.add_argument("--camera_idx", type=int, help="Index of which video source to use.", default=1)
    parser.add_argument("--num_detections", type=int, help="Number of detections to return.", default=3)
    parser.add_argument("--min_area_ratio", type=float, help="segment centroid min area ratio.", default=0.05)
    parser.add_argument("--batch_frames", type=int, default=16, help="frames per GPU pass (new)")
    parser.add_argument("--mode", choices=["contours", "labels"], default="contours", help="object definition (new)")
    parser.add_argument("--engine", choices=["synthetic", "torchvision", "torchvision-fp16"], default="synthetic",
                        help="class-score provider standing in for the Edge TPU model (new)")
    parser.add_argument("--contour_stage", choices=["gpu", "host", "cv2", "cv2-processes"], default="gpu",
                        help="where findContours / contourArea / moments run: the library's border following on host threads, "
                             "contour integrals on the GPU (default), OpenCV in this process, OpenCV in worker processes (new)")
    parser.add_argument("--port", default="[::]:50051")
    parser.set_defaults(keep_aspect_ratio=True)      # as the reference: always true (:246-259)
    return parser


def serve(argv=None):
    import cv2
    import grpc
    args = build_arg_parser().parse_args(argv)
    print("Loading {} with {} labels.".format(args.model, args.labels))
    engine = SyntheticEngine() if args.engine == "synthetic" else TorchvisionEngine(half=args.engine.endswith("fp16"))
    labels = hotpath.load_labels(args.labels)
    cap = cv2.VideoCapture(args.camera_idx)
    camera_res = int(cap.get(cv2.CAP_PROP_FRAME_WIDTH)), int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT))
    cap.release()
    with concurrent.futures.ThreadPoolExecutor(max_workers=10) as executor:
        stage = {"host": "host", "gpu": "gpu", "cv2": 0, "cv2-processes": CONTOUR_WORKERS}[args.contour_stage]
        future = executor.submit(recognize_and_segment, args.keep_aspect_ratio, args.camera_idx, engine, labels,
                                 args.min_area_ratio, args.batch_frames, args.mode, 0, None, stage)
        server = grpc.server(executor)
        add_SemanticSegmentationServicer_to_server(SemanticSegmentationServicer(args.num_detections, camera_res), server)
        server.add_insecure_port(args.port)
        server.start()
        future.result()
        server.stop(None)


if __name__ == "__main__":
    serve()
