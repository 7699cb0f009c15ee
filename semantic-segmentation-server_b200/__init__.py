"""B200-native hot path of goruck/semantic-segmentation-server: frame pre-process, logits -> label
map, per-label segment statistics.  Host code in Python, kernels in libssb200.so (C-ABI, CUDA
sm_100a).  Importing this package loads the library and fails if it is not built."""
from . import _cabi, contour_pool, hotpath, scheduler  # noqa: F401
from .hotpath import (Centroid, FrameGeometry, HostPipeline, HostStaging, IngestPipeline, SegmentContour, argmax_stats,  # noqa: F401
                      argmax_stats_nchw, blob_regions, candidates_from_moments_batch, contour_candidates_batch, contour_moments,
                      fold_region_histograms, fold_regions_batch, frames_to_device,
                      objects_from_regions,
                      create_pascal_label_colormap, foreground_label_set, foreground_mask, load_labels,
                      nearest_index_table, normalise, objects_from_contours, objects_from_stats,
                      preprocess_frames)

from .scheduler import FrameScheduler, aligned_chunks, bind_process_to_device_numa, partition_frames  # noqa: F401

__version__ = "0.1.0"
