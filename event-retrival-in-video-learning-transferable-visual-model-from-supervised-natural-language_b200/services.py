"""Host-side mirror of the reference's service layer for the retrieval hot path.

`EmbeddingService` keeps the reference's names, positional orders, defaults, return types and error behaviour
(Backend/services/embedding_service.py:70-536) but every arithmetic step runs in libevr.so on the B200:
  encode_image                    -> vit.VisionEncoder (through the `clip` face, clip.py)
  load + L2 normalise (:208-210)  -> search.FrameStore.from_numpy  (evr_store_prepare, HBM-resident fp16 store)
  np.dot + argsort (:314-320)     -> FrameStore.search             (evr_topk_search)
  np.dot(emb[i:i+1], q.T) (:274)  -> FrameStore.rescore            (evr_rescore)
The constructor takes the reference's own CacheService / PathService / DataService objects unchanged (they are
stdlib-only); `PathService`, `CacheService`, `DataService` below are small mirrors of exactly the methods this
path calls (path_service.py:47-93, cache_service.py:23-175, data_service.py:24-55) for standalone use.
Quirks of the reference that are fixed rather than copied are listed in INTEGRATION.md.
"""
import json
import os
import threading
import time

import numpy as np
import torch

from collections import OrderedDict, deque
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

from . import clip as _clip
from .preprocess import GpuPreprocess
from .search import FrameStore, shard_bounds

IMAGE_EXT = ('.jpg', '.jpeg', '.png')  # embedding_service.py:452


class PathService:
    """path_service.py:6-111 -- video_mapping.json lookup; `None` / unknown video -> first mapping entry."""

    def __init__(self, base_dir, metadata_dir, embedding_dir):
        self.base_dir, self.metadata_dir, self.embedding_dir = base_dir, metadata_dir, embedding_dir
        self.video_data_mapping = self._load_video_data_mapping()

    def _mapping_file(self):
        return os.path.join(self.metadata_dir, "video_mapping.json")

    def _load_video_data_mapping(self):
        if os.path.exists(self._mapping_file()):
            with open(self._mapping_file(), 'r', encoding='utf-8') as f:
                return json.load(f)
        return {}

    def save_video_data_mapping(self):
        with open(self._mapping_file(), 'w', encoding='utf-8') as f:
            json.dump(self.video_data_mapping, f, ensure_ascii=False, indent=4)

    def _entry(self, video_name, field, fallback):
        m = self.video_data_mapping
        if video_name and video_name in m:
            return os.path.normpath(m[video_name][field])
        if m:
            return m[next(iter(m))][field]
        return fallback

    def get_metadata_path(self, video_name=None):
        return self._entry(video_name, "metadata_file", os.path.join(self.metadata_dir, "output_samples.json"))

    def get_embeddings_path(self, video_name=None):
        return self._entry(video_name, "embeddings_file", os.path.join(self.embedding_dir, "image_embeddings.npy"))

    def add_video_mapping(self, video_name, metadata_file, embeddings_file, video_path, frames_dir):
        self.video_data_mapping[video_name] = {"metadata_file": metadata_file, "embeddings_file": embeddings_file,
                                               "video_path": video_path, "frames_dir": frames_dir}


class CacheService:
    """cache_service.py:7-241 -- keyed in-memory caches with a TTL (default 3600 s); same key construction."""

    def __init__(self, default_ttl=3600):
        self.default_ttl = default_ttl
        self._data = {}
        self._lock = threading.Lock()

    def _get(self, key):
        with self._lock:
            hit = self._data.get(key)
            if hit is None:
                return None
            value, created, ttl = hit
            if time.time() - created > ttl:
                del self._data[key]
                return None
            return value

    def _set(self, key, value, ttl=None):
        with self._lock:
            self._data[key] = (value, time.time(), self.default_ttl if ttl is None else ttl)

    def get_path_cache(self, video_name):
        return self._get(("paths", f"paths_{video_name or 'default'}"))

    def set_path_cache(self, video_name, embeddings_path, json_path, ttl=None):
        self._set(("paths", f"paths_{video_name or 'default'}"),
                  {'embeddings_path': embeddings_path, 'json_path': json_path}, ttl)

    def get_text_features(self, query, video_name):
        return self._get(("text", f"{query}_{video_name or 'default'}"))

    def set_text_features(self, query, video_name, features, ttl=None):
        self._set(("text", f"{query}_{video_name or 'default'}"), features, ttl)

    def get_embeddings(self, embeddings_path):
        return self._get(("emb", embeddings_path))

    def set_embeddings(self, embeddings_path, embeddings, ttl=None):
        self._set(("emb", embeddings_path), embeddings, ttl)

    def get_frames_list(self, json_path):
        return self._get(("frames", json_path))

    def set_frames_list(self, json_path, frames_list, ttl=None):
        self._set(("frames", json_path), frames_list, ttl)

    def get_search_results(self, query, video_name):
        return self._get(("search", f"{query}_{video_name or 'default'}"))

    def set_search_results(self, query, video_name, results, ttl=None):
        self._set(("search", f"{query}_{video_name or 'default'}"), results, ttl)

    def clear_all_caches(self):
        with self._lock:
            self._data.clear()


class DataService:
    """data_service.py:24-55 -- row i of the store <-> basename(metadata[i]["filepath"])."""

    def __init__(self, path_service, cache_service):
        self.path_service, self.cache_service = path_service, cache_service
        self._fps = {}
        self.frames_mapping = self.load_frames_mapping_from_json()   # data_service.py:22

    def load_frames_from_json(self, video_name=None):
        json_path = self.path_service.get_metadata_path(video_name)
        frames = self.cache_service.get_frames_list(json_path)
        if frames:
            return frames
        json_path = os.path.normpath(json_path)
        try:
            with open(json_path, 'r', encoding='utf-8') as f:
                samples = json.load(f)
            frames = [os.path.basename(s["filepath"]) for s in samples if "filepath" in s]
            self.cache_service.set_frames_list(json_path, frames)
            return frames
        except Exception as e:  # the reference swallows and returns []
            print(f"Error loading frames from JSON: {e}, path: {json_path}")
            return []

    def load_frames_mapping_from_json(self, video_name=None):
        """data_service.py:57-79 -- {frame file name: full path}; {} on any error."""
        json_path = os.path.normpath(self.path_service.get_metadata_path(video_name))
        try:
            with open(json_path, "r", encoding="utf-8") as f:
                return {Path(sample["filepath"]).name: sample["filepath"] for sample in json.load(f)}
        except Exception as e:
            print(f"Error loading frames mapping from JSON: {e}, path: {json_path}")
            return {}

    def load_json_data(self, video_name=None):
        """data_service.py:127-145 -- the metadata records of a video (file order); [] on any error."""
        json_path = self.path_service.get_metadata_path(video_name)
        try:
            with open(json_path, "r", encoding="utf-8") as f:
                return json.load(f)
        except Exception as e:
            print(f"Error loading data from JSON: {e}, path: {json_path}")
            return []

    def _video_fps(self, video_path):
        """fps used for `timestamp`: the file's own when it can be opened, else 25.0 (data_service.py:217-229).
        Probed once per (path, mtime) instead of on every event."""
        if not video_path or not os.path.exists(video_path):
            return 25.0
        key = (video_path, os.path.getmtime(video_path))
        fps = self._fps.get(key)
        if fps is None:
            fps = 25.0
            try:
                import cv2
                cap = cv2.VideoCapture(video_path)
                if cap.isOpened():
                    actual = cap.get(cv2.CAP_PROP_FPS)
                    if actual > 0:
                        fps = actual
                    cap.release()
            except Exception as e:
                print(f"Error getting video fps: {e}")
            self._fps[key] = fps
        return fps

    def get_video_duration(self, video_path):
        try:
            import cv2
            cap = cv2.VideoCapture(video_path)
            if not cap.isOpened():
                return 0
            fps, frames = cap.get(cv2.CAP_PROP_FPS), int(cap.get(cv2.CAP_PROP_FRAME_COUNT))
            cap.release()
            return round(frames / fps if fps > 0 else 0)
        except Exception:
            return 0

    def format_event_for_frontend(self, frame_data):
        """data_service.py:147-251 -- the event dict the UI receives.  `confidence` starts at 0.7, becomes the best OCR
        confidence when there is OCR text, the best object confidence when that is higher, and the CLIP similarity
        when that is higher still; `detection_type` names the winner."""
        video_path = frame_data.get('video', '')
        video_id = f"video-{Path(video_path).stem}" if video_path else 'unknown'
        category, description, detection_type = "Unknown", "Event detected", "unknown"
        confidence, text_confidence, object_confidence = 0.7, 0.0, 0.0

        texts = (frame_data.get('text_detections') or {}).get('detections') or []
        if texts:
            best = max(texts, key=lambda d: d.get('confidence', 0))
            label = best.get('label')
            category = label.split(' ')[0] if label else "Unknown"
            text_confidence = _pyfloat(best.get('confidence', 0.7))
            description = best.get('label', 'Event detected')
            detection_type, confidence = "text", text_confidence

        objects = (frame_data.get('object_detections') or {}).get('detections') or []
        if objects:
            best = max(objects, key=lambda d: d.get('confidence', 0))
            object_confidence = _pyfloat(best.get('confidence', 0.5))
            if object_confidence > text_confidence:
                category = best.get('label', 'Unknown')
                description = f"Object detected: {category}"
                detection_type, confidence = "object", object_confidence

        clip_similarity = frame_data.get('clip_similarity', 0.0)
        clip_similarity = 0.0 if clip_similarity is None else _pyfloat(clip_similarity)
        if clip_similarity > confidence:
            detection_type, confidence = "clip", clip_similarity

        frame_idx = frame_data.get('frameidx', 0)
        return {"id": f"event-{frame_idx}", "videoId": video_id, "title": f"Event at frame {frame_idx}",
                "description": description, "timestamp": _pyfloat(frame_idx / self._video_fps(video_path)),
                "duration": 5, "category": category, "confidence": confidence, "text_confidence": text_confidence,
                "object_confidence": object_confidence, "clip_similarity": clip_similarity,
                "detection_type": detection_type, "thumbnailUrl": frame_data.get('filepath')}


def _pyfloat(x):
    return float(x) if type(x).__module__ == "numpy" else x


def _dist_context(rank=None, world=None, barrier=None):
    """(rank, world, barrier) of a data-parallel call: explicit arguments win, else the initialised process group,
    else one process.  `barrier` is a callable every rank calls at the same points."""
    if rank is None or world is None:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            rank, world = dist.get_rank(), dist.get_world_size()
            barrier = barrier or dist.barrier
        else:
            rank, world = 0, 1
    if world > 1 and barrier is None:
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()):
            raise ValueError("world > 1 needs an initialised torch.distributed process group or an explicit barrier=")
        barrier = dist.barrier
    return int(rank), int(world), (barrier or (lambda: None))


def shard_rows_to_npy(path, n_rows, lo, rows, rank=0, world=1, barrier=None):
    """One `<video>_embeddings.npy` ([n_rows, D], the layout of embedding_service.py:505) written by `world` ranks,
    rank r holding rows [lo, lo + len(rows)).  One rank: `np.save`.  Several: rank 0 creates the file at its final
    size (`np.lib.format.open_memmap`: the same NPY v1 header `np.save` writes), a barrier, every rank writes its
    slice through a memory map, a barrier.  No row crosses a link: the exchange is the shared file system."""
    if world == 1:
        np.save(path, rows)
        return
    barrier = barrier or (lambda: None)
    mm = None
    if rank == 0:
        mm = np.lib.format.open_memmap(path, mode="w+", dtype=rows.dtype, shape=(int(n_rows), rows.shape[1]))
    barrier()
    if mm is None:
        mm = np.load(path, mmap_mode="r+")
        if mm.shape != (int(n_rows), rows.shape[1]) or mm.dtype != rows.dtype:
            raise RuntimeError(f"{path}: rank {rank} sees {mm.shape} {mm.dtype}, expected {(n_rows, rows.shape[1])} "
                               f"{rows.dtype}")
    if len(rows):
        mm[lo:lo + len(rows)] = rows
    # no msync: like np.save (which does not fsync either) the rows are in the page cache, where every process of the
    # node sees them; forcing them to the device cost 20-60 ms per 16-130 MB file with 2-8 ranks writing at once
    del mm
    barrier()


def encode_frames_sharded(encoder, host_batches, n_rows, lo, path, rank=None, world=None, barrier=None,
                          out_dtype=torch.float16):
    """The data-parallel ingest without its JPEG decode: this rank's contiguous slice of a frame list, already decoded
    (`host_batches`: pinned uint8 [B,S,S,3] or float [B,3,S,S] host tensors, global rows [lo, lo + sum(B))), goes
    through `VisionEncoder.encode_stream` (H2D copy of batch i+1 under the tower pass of batch i, embeddings back to
    pinned memory) and the unit-norm rows land in the one `.npy` of `n_rows` rows every rank writes
    (`shard_rows_to_npy`).  Returns this rank's rows.  What `extract_and_save_embeddings_from_folder` does after PIL."""
    rank, world, barrier = _dist_context(rank, world, barrier)
    rows = [e.numpy() for e in encoder.encode_stream(host_batches, normalise=True, out_dtype=out_dtype)]
    rows = np.concatenate(rows) if rows else np.zeros((0, encoder.out_dim),
                                                      np.float16 if out_dtype == torch.float16 else np.float32)
    shard_rows_to_npy(path, n_rows, lo, rows, rank, world, barrier)
    return rows


def decode_frames(paths, workers=None, window=None):
    """Ordered, bounded-lookahead JPEG decode on host threads: yields `np.uint8 [H, W, 3]` (or the exception that the
    decode raised) for every path, in order.  The reference decodes one frame at a time on the request thread
    (`Image.open(p).convert('RGB')`, embedding_service.py:475); PIL's decoder releases the GIL, so `workers` threads give
    the same pixels `workers` times faster, and at most `window` decoded frames wait for the GPU at any time."""
    from PIL import Image

    def one(path):
        try:
            return np.array(Image.open(path).convert('RGB'))
        except Exception as e:  # the caller turns this into a zero row (embedding_service.py:477-480)
            return e

    workers = min(16, os.cpu_count() or 1) if workers is None else int(workers)
    if workers <= 1:
        for path in paths:
            yield one(path)
        return
    window = window or 4 * workers
    with ThreadPoolExecutor(max_workers=workers) as pool:
        queue, it = deque(), iter(paths)
        for path in it:
            queue.append(pool.submit(one, path))
            if len(queue) >= window:
                break
        while queue:
            frame = queue.popleft().result()
            nxt = next(it, None)
            if nxt is not None:
                queue.append(pool.submit(one, nxt))
            yield frame


class EmbeddingService:
    """Drop-in for Backend/services/embedding_service.py:EmbeddingService (same constructor positionals).

    Extra keyword arguments (all optional, absent in the reference): `clip_module` (defaults to this package's
    `clip`), `model_name`, `load_kwargs` (forwarded to clip.load, e.g. state_dict= / random_init=),
    `text_encoder` (callable str -> [1, D] array, used while the text tower is not built: SURVEY 8f row 1),
    `checkpoint_path` (fine-tuned CLIPWithClassifier checkpoint, embedding_service.py:93), `gpu_preprocess` (ingest
    resizes/crops on the device, bit-identical to PIL; False keeps the per-frame host `preprocess`), `gpu_jpeg` (ingest
    decodes `.jpg` files with nvJPEG on the device instead of PIL on host threads; default off / $EVR_GPU_JPEG)."""

    def __init__(self, cache_service, path_service, data_service, device="cuda", clip_module=None,
                 model_name="ViT-B/32", load_kwargs=None, text_encoder=None, checkpoint_path=None,
                 gpu_preprocess=True, gpu_jpeg=None):
        self.cache_service, self.path_service, self.data_service = cache_service, path_service, data_service
        self.device = device
        self._clip = clip_module or _clip
        self._model_name = model_name
        self._load_kwargs = dict(load_kwargs or {})
        self.text_encoder = text_encoder
        self.gpu_preprocess = bool(gpu_preprocess)
        # JPEG decode on the device (nvJPEG; preprocess.JpegDecoder): opt-in, because its pixels are not bit-identical
        # to PIL's (embeddings agree to cosine >= 0.999); the default keeps the reference's own decoder
        self.gpu_jpeg = bool(int(os.environ.get("EVR_GPU_JPEG", "0"))) if gpu_jpeg is None else bool(gpu_jpeg)
        self.decode_workers = None  # host threads decoding JPEGs during ingest (None: min(16, cores); 1: serial)
        self.original_model, self.preprocess = self._clip.load(model_name, device=device, **self._load_kwargs)
        self.finetuned_model = None
        self.active_model = "original"
        self.checkpoint_path = checkpoint_path
        self._stores = {}              # embeddings_path -> (mtime, FrameStore, source dtype)
        self._scan_scores = OrderedDict()  # (model, query, video) -> (FrameStore, {frame name: cosine}) of past scans
        self._lock = threading.RLock()  # the Flask dev server is threaded (app.py:968); the reference holds no locks
        if checkpoint_path and os.path.exists(checkpoint_path):
            try:
                self._load_finetuned_model(checkpoint_path)
            except Exception as e:
                print(f"Error loading fine-tuned model: {e}")

    # ---- models (a6) ------------------------------------------------------------------------------
    def _load_finetuned_model(self, checkpoint_path):
        """embedding_service.py:103-120: CLIPWithClassifier checkpoint, CLIP weights under `clip_model.`; at serve
        time only encode_image + L2 normalise of that wrapper is used (:36-49), the classifier head is not."""
        model, _ = self._clip.load(checkpoint_path, device=self.device)
        self.finetuned_model = _FinetunedImageEncoder(model)

    def set_active_model(self, model_name):
        if model_name == "original":
            self.active_model = "original"
            return True
        if model_name == "finetuned":
            if self.finetuned_model is not None:
                self.active_model = "finetuned"
                return True
            print("Fine-tuned model is not loaded!")
            return False
        print(f"Unknown model name: {model_name}")
        return False

    def get_active_model_name(self):
        return self.active_model

    def _encode_images(self, batch, model_key=None):
        """-> numpy [B, D] un-normalised, in the model dtype (original) / unit-norm fp32 (fine-tuned wrapper)."""
        key = model_key or self.active_model
        with torch.no_grad():
            if key == "finetuned" and self.finetuned_model is not None:
                return self.finetuned_model(batch).cpu().numpy()
            return self.original_model.encode_image(batch).cpu().numpy()

    # ---- text features (a8) -----------------------------------------------------------------------
    def get_text_features(self, query, video_name=None):
        cache_key = f"{self.active_model}_{query}_{video_name or 'default'}"
        features = self.cache_service.get_text_features(cache_key, video_name)
        if features is not None:
            return features
        if self.text_encoder is not None:
            features = np.asarray(self.text_encoder(query), dtype=np.float32).reshape(1, -1)
        else:
            model = self.original_model
            if self.active_model == "finetuned" and self.finetuned_model is not None:
                model = self.finetuned_model.clip_model
            text_input = self._clip.tokenize([query]).to(self.device)
            with torch.no_grad():
                features = model.encode_text(text_input).cpu().numpy()
        features = features / np.linalg.norm(features, axis=-1, keepdims=True)
        self.cache_service.set_text_features(cache_key, video_name, features)
        return features

    # ---- store (a7) -------------------------------------------------------------------------------
    def _frame_store(self, video_name=None):
        """HBM-resident FrameStore of the video's `.npy` (re-read when the file changes); None if missing."""
        path = self.path_service.get_embeddings_path(video_name)
        if not os.path.exists(path):
            print(f"Warning: Embeddings file not found: {path}")
            return None, path
        mtime = os.path.getmtime(path)
        with self._lock:
            hit = self._stores.get(path)
            if hit is not None and hit[0] == mtime:
                return hit, path
            raw = np.load(path)
            entry = (mtime, FrameStore.from_numpy(raw, device=self.device), raw.dtype)
            self._stores[path] = entry
            return entry, path

    def get_embeddings(self, video_name=None):
        """Unit-norm rows as a numpy array of the file's dtype (embedding_service.py:186-217).  The rows are the
        device store read back (normalised by evr_store_prepare, fp16 in HBM), not a host re-computation."""
        try:
            entry, path = self._frame_store(video_name)
            if entry is None:
                return None
            cached = self.cache_service.get_embeddings(path)
            if cached is not None and len(cached) == entry[1].N:
                return cached
            emb = entry[1].rows.cpu().numpy().astype(entry[2] if entry[2] in (np.float16, np.float32)
                                                     else np.float32, copy=False)
            self.cache_service.set_embeddings(path, emb)
            return emb
        except Exception as e:
            print(f"Error loading embeddings: {e}")
            return None

    def _frames(self, video_name):
        json_path = self.path_service.get_metadata_path(video_name)
        frames = self.cache_service.get_frames_list(json_path)
        if frames is None:
            frames = self.data_service.load_frames_from_json(video_name)
            self.cache_service.set_frames_list(json_path, frames)
        return frames

    # ---- re-score (a11) ---------------------------------------------------------------------------
    def extract_query_confidence(self, frame_path, query, video_name=None):
        """cosine(frame row, text features) as a python float; 0.0 on a miss or any error (:219-282).

        The strategies call this once per candidate right after `search_top_frames(query, top_k*3, video)`
        (query_strategies.py:55, :100): the scan has already produced exactly this dot product, so it is answered from
        the scan's own (frame, score) pairs (SURVEY.md 8g "search_top_frames drops scores") and only a frame the scan
        did not return costs a kernel (`evr_rescore`)."""
        try:
            text_features = self.get_text_features(query, video_name)
            entry, _ = self._frame_store(video_name)
            if entry is None:
                return 0.0
            known = self._scan_scores.get((self.active_model, query, video_name or 'default'))
            if known is not None and known[0] is entry[1]:
                score = known[1].get(frame_path)
                if score is None:
                    score = known[1].get(os.path.basename(frame_path))
                if score is not None:
                    return score
            index = self._row_index(video_name).get(frame_path)
            if index is None:
                index = self._row_index(video_name).get(os.path.basename(frame_path))
            if index is None or index >= entry[1].N:
                return 0.0
            q = torch.from_numpy(np.ascontiguousarray(text_features, dtype=np.float32).reshape(-1))
            return float(entry[1].rescore(q, [index]).item())
        except Exception as e:
            print(f"Error extracting query confidence: {e}")
            return 0.0

    def _row_index(self, video_name):
        """frame name -> first row with that name: what `frames.index(name)` (:264) finds, without its O(N) walk."""
        json_path = self.path_service.get_metadata_path(video_name)
        frames = self._frames(video_name)
        with self._lock:
            hit = getattr(self, "_row_maps", {}).get(json_path)
            if hit is not None and hit[0] is frames:
                return hit[1]
            table = {}
            for i, f in enumerate(frames):
                table.setdefault(f, i)
                table.setdefault(os.path.basename(f), i)
            self._row_maps = dict(getattr(self, "_row_maps", {}))
            self._row_maps[json_path] = (frames, table)
            return table

    # ---- the scan (a9, a10) -----------------------------------------------------------------------
    def _scan(self, features, top_k, video_name=None):
        """-> [(frame name, cosine, store row)] best first: the fused scan (evr_topk_search) over the video's store."""
        entry, _ = self._frame_store(video_name)
        if entry is None:
            return []
        store = entry[1]
        q = torch.from_numpy(np.ascontiguousarray(np.asarray(features, dtype=np.float32).reshape(1, -1)))
        scores, idx = store.search(q, min(int(top_k), store.N))
        scores, idx = scores[0].cpu().numpy(), idx[0].cpu().numpy()
        frames = self._frames(video_name)
        return [(frames[int(i)], float(s), int(i)) for i, s in zip(idx, scores)]

    def search_with_scores(self, features, top_k, video_name=None):
        """[(frame name, cosine)] in descending score: the pairs the reference builds at :330-333 before it drops
        the scores.  `features` [1, D] / [D] unit-norm query."""
        return [(f, s) for f, s, _ in self._scan(features, top_k, video_name)]

    def search_top_frames(self, query, top_k, video_name=None):
        """list of frame file names, best first; [] on any error; memoised like the reference (:284-344)."""
        try:
            cache_key = f"search_{self.active_model}_{query}_{top_k}"
            cached = self.cache_service.get_search_results(cache_key, video_name)
            if cached is not None:
                return cached
            text_features = self.get_text_features(query, video_name)
            hits = self._scan(text_features, top_k, video_name)
            self._remember_scores(query, video_name, hits)
            top_frames = [f for f, _, _ in hits]
            if top_frames or self._frame_store(video_name)[0] is not None:
                self.cache_service.set_search_results(cache_key, video_name, top_frames)
            return top_frames
        except Exception as e:
            print(f"Error in search_top_frames: {e}")
            return []

    def _remember_scores(self, query, video_name, hits, keep=256):
        """(model, query, video) -> (store, {frame: cosine}) for the candidates of the last scans.  Only the first row
        carrying a frame name is recorded, because that is the row `frames.index(name)` scores (:264).  Bounded: the
        oldest of `keep` queries is dropped."""
        entry, _ = self._frame_store(video_name)
        if entry is None or not hits:
            return
        first_row = self._row_index(video_name)
        key = (self.active_model, query, video_name or 'default')
        with self._lock:
            old = self._scan_scores.get(key)
            table = dict(old[1]) if old is not None and old[0] is entry[1] else {}
            for name, score, row in hits:
                if first_row.get(name) == row:
                    table[name] = score
                    table.setdefault(os.path.basename(name), score)
            self._scan_scores[key] = (entry[1], table)
            self._scan_scores.move_to_end(key)
            while len(self._scan_scores) > keep:
                self._scan_scores.popitem(last=False)

    def search_top_frames_by_image(self, image_features, top_k, video_name=None):
        """Same scan with a given [1, D] query vector, no result cache (:346-392)."""
        try:
            return [f for f, _ in self.search_with_scores(image_features, top_k, video_name)]
        except Exception as e:
            print(f"Error in search_top_frames_by_image: {e}")
            return []

    # ---- encode (a2, a5) --------------------------------------------------------------------------
    def extract_image_embedding(self, image_path):
        """[1, D] unit-norm numpy embedding of one image file, None on error (:394-423)."""
        try:
            from PIL import Image
            image = self.preprocess(Image.open(image_path)).unsqueeze(0).to(self.device)
            emb = self._encode_images(image)
            return emb / np.linalg.norm(emb, axis=-1, keepdims=True)
        except Exception as e:
            print(f"Error extracting embedding from {image_path}: {e}")
            return None

    def extract_and_save_embeddings_from_folder(self, folder_path, model_name=None, video_name=None, rank=None,
                                                world=None, barrier=None):
        """Encode every frame of a folder (sorted file names) and write `<video>_embeddings.npy` [N, D] unit-norm
        (:425-536).  Returns the path, or None when the folder has no images.  Differences from the reference, all
        from SURVEY.md 8g: the model is chosen per call without flipping `active_model` under other threads; an
        unknown video gets `<embedding_dir>/<video>_embeddings.npy` instead of overwriting the first video's file;
        an undecodable frame becomes a zero row (not a NameError / NaN row).

        Data-parallel ingest (SURVEY.md 8e "Encode"; absent from the reference, which is single-device): when
        `torch.distributed` is initialised (or `rank` / `world` are given) every rank calls this with the same
        arguments, the sorted frame list is cut into contiguous rank slices (`search.shard_bounds`), each rank encodes
        its slice and writes its rows into one pre-sized `.npy` (`shard_rows_to_npy`), and rank 0 stamps the metadata.
        Frames are independent, so there is no collective on the data path; row i of the file is bit for bit the row
        a single GPU writes (the tower is batch-invariant, the final normalisation is per row)."""
        key = self.active_model
        if model_name:
            key = "original" if (model_name == "finetuned" and self.finetuned_model is None) else model_name
            if key not in ("original", "finetuned"):
                print(f"Unknown model name: {model_name}")
                key = self.active_model
        mapping = getattr(self.path_service, "video_data_mapping", {})
        if video_name and video_name not in mapping and getattr(self.path_service, "embedding_dir", None):
            embeddings_path = os.path.join(self.path_service.embedding_dir, f"{video_name}_embeddings.npy")
        else:
            embeddings_path = self.path_service.get_embeddings_path(video_name)
        os.makedirs(os.path.dirname(embeddings_path) or ".", exist_ok=True)
        frame_files = sorted(f for f in os.listdir(folder_path) if f.endswith(IMAGE_EXT))
        if not frame_files:
            print(f"No image files found in {folder_path}")
            return None
        rank, world, barrier = _dist_context(rank, world, barrier)
        lo, hi = shard_bounds(len(frame_files), world, rank)
        embeddings, bad_rows = self._encode_frames(folder_path, frame_files[lo:hi], key)
        norm = np.linalg.norm(embeddings, axis=-1, keepdims=True)
        embeddings = embeddings / np.where(norm > 1e-8, norm, 1).astype(embeddings.dtype)
        embeddings[bad_rows] = 0
        shard_rows_to_npy(embeddings_path, len(frame_files), lo, embeddings, rank, world, barrier)
        if rank == 0:
            self._stamp_metadata(video_name, key)
        return embeddings_path

    def _embedding_dtype(self, key):
        if key == "finetuned" and self.finetuned_model is not None:
            return np.float32
        return np.float16 if getattr(self.original_model, "dtype", torch.float32) == torch.float16 else np.float32

    def _encode_frames(self, folder_path, frame_files, key):
        """-> ([n, D] un-normalised rows of these files in the model dtype, indices of undecodable frames)."""
        from PIL import Image
        S = self.original_model.visual.input_resolution
        all_embeddings, bad_rows = [], []
        if not frame_files:   # more ranks than frames
            return np.zeros((0, self.original_model.visual.output_dim), self._embedding_dtype(key)), bad_rows
        if getattr(self.preprocess, "evr_n_px", None) == S and self.gpu_preprocess and self.gpu_jpeg:
            return self._encode_frames_nvjpeg(folder_path, frame_files, key, S)
        if getattr(self.preprocess, "evr_n_px", None) == S and self.gpu_preprocess:
            # device path: PIL only decodes; Resize(bicubic) + CenterCrop run on the GPU bit-identically to PIL
            # (preprocess.py) and ToTensor + Normalize are fused into the encoder's im2col.  Frames of one size are
            # batched together; the row order stays the sorted file order.
            pre = self._gpu_pre = getattr(self, "_gpu_pre", None) or GpuPreprocess(S, self.device)
            shape, stage, stage_np, count = None, None, None, 0
            stages = self._ingest_stages = getattr(self, "_ingest_stages", None) or {}  # pinned once, kept across calls

            def flush():
                nonlocal count
                if count:
                    u8 = pre(stage[:count])  # pinned -> device, resize + crop on the GPU
                    all_embeddings.append(self._encode_images(u8, key))  # returns host rows: the stage is free again
                    count = 0

            paths = [os.path.join(folder_path, name) for name in frame_files]
            for i, frame in enumerate(decode_frames(paths, self.decode_workers)):
                if isinstance(frame, Exception):
                    print(f"Error preprocessing image {frame_files[i]}: {frame}")
                    frame = np.zeros(shape or (S, S, 3), dtype=np.uint8)
                    bad_rows.append(i)
                if frame.shape != shape or count == stage.shape[0]:
                    flush()
                    if frame.shape != shape:  # frames of one size share a pinned staging batch (one per ingest, usually)
                        shape = frame.shape
                        stage = stages.get(shape)
                        if stage is None:
                            if len(stages) >= 2:
                                stages.clear()
                            # 64 frames per staging batch: page-locking costs ~1 ms per MB, a 720p batch is 177 MB
                            stage = stages[shape] = torch.empty((64,) + shape, dtype=torch.uint8).pin_memory()
                        stage_np = stage.numpy()
                stage_np[count] = frame  # numpy memcpy (releases the GIL; torch's copy_ would spin up its thread pool)
                count += 1
            flush()
        else:
            batch_size = 32  # :461
            for i in range(0, len(frame_files), batch_size):
                batch = []
                for j, name in enumerate(frame_files[i:i + batch_size]):
                    try:
                        batch.append(self.preprocess(Image.open(os.path.join(folder_path, name)).convert('RGB')))
                    except Exception as e:
                        print(f"Error preprocessing image {name}: {e}")
                        batch.append(torch.zeros(3, S, S))
                        bad_rows.append(i + j)
                all_embeddings.append(self._encode_images(torch.stack(batch).to(self.device), key))
        return np.vstack(all_embeddings), bad_rows

    def _encode_frames_nvjpeg(self, folder_path, frame_files, key, S, batch=256, workers=3):
        """The ingest loop with the JPEG decode on the device.  The sorted frame list is cut into runs of `batch` files;
        `workers` host threads, each with its own nvJPEG state and CUDA stream, read a run, split it into segments of
        same-size baseline JPEGs (sizes from the SOF headers) and decode each segment with one batched call, ahead of
        the consumer.  nvJPEG's batched decode costs ~20 ms per call almost independently of the batch (its GPU Huffman
        stage is a few long sequential chains per image), so the batches are large (256 frames of 720p = 0.7 GB of RGB)
        and several are in flight.  The calling thread takes the decoded segments in order: resize + crop
        (`GpuPreprocess`), tower, rows back.  A file nvJPEG cannot take (PNG, progressive or damaged JPEG) goes through
        PIL for that frame; an undecodable frame is a zero row, as in the PIL path."""
        from PIL import Image
        from ._lib import EvrError
        from .preprocess import JpegDecoder, jpeg_frame_size
        pre = self._gpu_pre = getattr(self, "_gpu_pre", None) or GpuPreprocess(S, self.device)
        self._jpeg = getattr(self, "_jpeg", None) or JpegDecoder(self.device)
        pool_state = self._jpeg_workers = getattr(self, "_jpeg_workers", None) or threading.local()
        rows = [None] * len(frame_files)
        bad_rows = []
        dev = torch.device(self.device)

        def via_pil(i):
            try:
                frame = np.array(Image.open(os.path.join(folder_path, frame_files[i])).convert('RGB'))
                rows[i] = self._encode_images(pre(frame), key)[0]
            except Exception as e:
                print(f"Error preprocessing image {frame_files[i]}: {e}")
                bad_rows.append(i)

        def decode_run(idx):
            """worker thread: files of one run -> [(indices, device frames | None, event | None)] in file order"""
            if not hasattr(pool_state, "dec"):
                pool_state.dec = JpegDecoder(dev)
                pool_state.stream = torch.cuda.Stream(device=dev)
            segments = []           # [indices, size, blobs]
            for i in idx:
                blob, size = None, None
                if frame_files[i].lower().endswith(('.jpg', '.jpeg')):
                    try:
                        with open(os.path.join(folder_path, frame_files[i]), "rb") as f:
                            blob = f.read()
                        size = jpeg_frame_size(blob)     # None: not a baseline JPEG
                    except OSError:
                        size = None
                if size is not None and segments and segments[-1][1] == size:
                    segments[-1][0].append(i)
                    segments[-1][2].append(blob)
                else:
                    segments.append([[i], size, [blob]])
            out = []
            for seg_idx, size, blobs in segments:
                if size is None:
                    out.append((seg_idx, None, None))
                    continue
                try:
                    with torch.cuda.device(dev), torch.cuda.stream(pool_state.stream):
                        frames = pool_state.dec.decode(blobs, size)
                        done = torch.cuda.Event()
                        done.record(pool_state.stream)
                    out.append((seg_idx, frames, done))
                except EvrError:
                    out.append((seg_idx, None, None))
            return out

        runs = [list(range(i, min(i + batch, len(frame_files)))) for i in range(0, len(frame_files), batch)]
        with ThreadPoolExecutor(max_workers=workers) as pool:
            pending, nxt = deque(), 0
            while nxt < len(runs) or pending:
                while nxt < len(runs) and len(pending) < workers:
                    pending.append(pool.submit(decode_run, runs[nxt]))
                    nxt += 1
                for idx, frames, done in pending.popleft().result():
                    if frames is None:     # not a baseline JPEG (or one of the segment is not): frame by frame via PIL
                        for i in idx:
                            via_pil(i)
                        continue
                    torch.cuda.current_stream(dev).wait_event(done)
                    frames.record_stream(torch.cuda.current_stream(dev))
                    emb = self._encode_images(pre(frames), key)
                    for i, row in zip(idx, emb):
                        rows[i] = row
                    del frames
        D = self.original_model.visual.output_dim
        out = np.zeros((len(frame_files), D), self._embedding_dtype(key))
        for i, row in enumerate(rows):
            if row is not None:
                out[i] = row
        return out, sorted(bad_rows)

    def _stamp_metadata(self, video_name, model_key):
        """:509-526 -- record which model produced the store in the video's metadata JSON."""
        metadata_path = self.path_service.get_metadata_path(video_name)
        mapping = getattr(self.path_service, "video_data_mapping", {})
        if (video_name and video_name not in mapping) or not os.path.exists(metadata_path):
            return
        try:
            with open(metadata_path, 'r', encoding='utf-8') as f:
                metadata = json.load(f)
            if isinstance(metadata, list):
                for item in metadata:
                    item['embedding_model'] = model_key
            elif isinstance(metadata, dict):
                metadata['embedding_model'] = model_key
            with open(metadata_path, 'w', encoding='utf-8') as f:
                json.dump(metadata, f, ensure_ascii=False, indent=2)
        except Exception as e:
            print(f"Error updating metadata with model info: {e}")


class _FinetunedImageEncoder:
    """Serve-time behaviour of CLIPWithClassifier.forward(images) (embedding_service.py:36-49): encode_image in
    fp32, then L2 normalise.  (The classifier head is never evaluated on this path.)"""

    def __init__(self, clip_model):
        self.clip_model = clip_model.float()

    def __call__(self, images, texts=None, get_embeddings=False):
        # uint8 [B,S,S,3] frames (device preprocessing path) go straight in: ToTensor + Normalize are fused in the tower
        f = self.clip_model.encode_image(images if images.dtype == torch.uint8 else images.float()).float()
        return f / f.norm(dim=-1, keepdim=True)

    def eval(self):
        return self


class SearchService:
    """Drop-in for the part of Backend/services/search_service.py that `app.py` reaches on this path (SURVEY.md 8a
    row a15): same constructor positionals, `search_by_image(image_url, adaptive_threshold, top_k, video_name=None,
    preprocess=None)` (:611-705), `search_semantic_with_clip` (:121-177), `search_frames_by_keyword` (:25-58, the
    callback the keyword strategies take), `sort_by_confidence` (:605-609).  The other `search_by_*` methods of the
    reference class duplicate query_strategies.py and are not called by `app.py`; they forward
    to the strategy functions here."""

    def __init__(self, embedding_service, data_service, path_service, cache_service):
        self.embedding_service, self.data_service = embedding_service, data_service
        self.path_service, self.cache_service = path_service, cache_service

    # ---- OCR keyword lookup (metadata only; no arithmetic) ----------------------------------------
    def search_frames_by_keyword(self, keyword, top_k, video_name=None):
        """frame ids whose OCR text contains the keyword (accent-insensitive), ranked by the confidence of the FIRST
        matching detection of each frame (the reference `break`s at the first hit, :49-55), best `top_k`."""
        from ._unaccent import unidecode
        needle = unidecode(keyword.lower())
        matches = []
        for record in self.data_service.load_json_data(video_name):
            for det in record.get("text_detections", {}).get("detections", []):
                label = det.get("label", "")
                if label and needle in unidecode(label.lower()):
                    matches.append((record.get("frameid", ""), det.get("confidence", 0)))
                    break
        matches.sort(key=lambda m: m[1], reverse=True)
        return [frameid for frameid, _ in matches[:top_k]]

    # ---- CLIP text search with a threshold ---------------------------------------------------------
    def search_semantic_with_clip(self, query, adaptive_threshold, top_k, video_name=None):
        """:121-177 -- `query_by_text_with_adaptive_threshold` without the text preprocessing step, over this
        service's own embedding / data services."""
        try:
            svc = self.embedding_service
            by_frameidx = {}
            for item in self.data_service.load_json_data(video_name):
                by_frameidx.setdefault(item.get('frameidx'), item)
            results = []
            for frame_name in svc.search_top_frames(query, top_k * 3, video_name):
                try:
                    record = by_frameidx.get(int(Path(frame_name).stem))
                    if not record:
                        continue
                    confidence = svc.extract_query_confidence(frame_name, query, video_name)
                    if confidence >= adaptive_threshold:
                        event = self.data_service.format_event_for_frontend(dict(record, clip_similarity=confidence))
                        event["clip_similarity"] = confidence
                        results.append(event)
                except Exception as e:
                    print(f"Error processing frame {frame_name}: {e}")
            results.sort(key=lambda x: x.get("clip_similarity", 0), reverse=True)
            return results[:top_k]
        except Exception as e:
            print(f"Error in semantic search: {e}")
            return []

    # ---- image query -------------------------------------------------------------------------------
    @staticmethod
    def _open_image(image_url):
        import base64
        from io import BytesIO
        from PIL import Image
        if hasattr(image_url, "convert"):
            return image_url.convert('RGB')
        if image_url.startswith(('http://', 'https://')):
            import requests  # the reference downloads the URL (:633-636); offline boxes raise here -> []
            return Image.open(BytesIO(requests.get(image_url, timeout=10).content)).convert('RGB')
        if image_url.startswith('data:image'):
            # the reference hands the base64 TEXT to PIL (:639-640), which can never decode; the payload is decoded here
            return Image.open(BytesIO(base64.b64decode(image_url.split(',')[1]))).convert('RGB')
        return Image.open(image_url).convert('RGB')

    def search_by_image(self, image_url, adaptive_threshold, top_k, video_name=None, preprocess=None,
                        format_event_for_frontend=None):
        """:611-705 -- encode the query image, scan with `top_k * 3`, keep candidates whose cosine passes
        `adaptive_threshold`, best `top_k` events (`confidence` = `clip_similarity` = the cosine).

        Difference from the reference: a candidate's similarity is the scan's own score -- the dot
        product of its store row with the query -- instead of re-embedding the frame through
        `extract_image_embedding(<basename>)` (:679), which fails outside the frames directory and silently drops the
        candidate, and costs up to 3 top_k extra tower passes when it does not."""
        try:
            svc = self.embedding_service
            fmt = format_event_for_frontend or self.data_service.format_event_for_frontend
            preprocess = preprocess if preprocess is not None else svc.preprocess
            image = preprocess(self._open_image(image_url)).unsqueeze(0).to(svc.device)
            with torch.no_grad():
                if svc.get_active_model_name() == "finetuned" and svc.finetuned_model is not None:
                    feats = svc.finetuned_model(image)
                else:
                    feats = svc.original_model.encode_image(image)
                feats = feats.float()
                feats = (feats / feats.norm(dim=-1, keepdim=True)).cpu().numpy()
            by_frameidx = {}
            for item in self.data_service.load_json_data(video_name):
                by_frameidx.setdefault(item.get('frameidx'), item)
            results = []
            for frame_name, similarity in svc.search_with_scores(feats, top_k * 3, video_name):
                try:
                    record = by_frameidx.get(int(Path(frame_name).stem))
                    if record and similarity >= adaptive_threshold:
                        event = fmt(dict(record, clip_similarity=float(similarity)))
                        event["clip_similarity"] = float(similarity)
                        event["confidence"] = float(similarity)
                        results.append(event)
                except Exception as e:
                    print(f"Error processing frame {frame_name}: {e}")
            results.sort(key=lambda x: x.get("clip_similarity", 0), reverse=True)
            return results[:top_k]
        except Exception as e:
            print(f"Error in image search: {e}")
            return []

    # ---- the reference class's duplicates of the strategies ---------------------------------------
    def _strategy_args(self):
        return dict(video_data_mapping=getattr(self.path_service, "video_data_mapping", {}))

    def search_by_keyword(self, query, adaptive_threshold, top_k, video_name=None):
        from . import query_strategies as qs
        return qs.query_by_keyword(query, adaptive_threshold, top_k, self.search_frames_by_keyword,
                                   self.data_service.format_event_for_frontend, video_name=video_name,
                                   **self._strategy_args())

    def search_by_object(self, query, adaptive_threshold, top_k, video_name=None):
        from . import query_strategies as qs
        return qs.query_by_object(query, adaptive_threshold, top_k, self.data_service.format_event_for_frontend,
                                  video_name=video_name, **self._strategy_args())

    def sort_by_confidence(self, results):
        return sorted(results, key=lambda x: x.get("confidence", 0), reverse=True) if results else []


def frame_metadata_items(folder_path, output_json_path, path_video):
    """Detector-free stand-in for `process_images_in_folder(folder_path, output_json_path, path_video)`
    (JSON_sample_DOC.py:90): the same item layout (:74-84) with empty OCR / object lists -- EasyOCR and YOLO are outside
    this path -- written in sorted frame order with the ingest's own (case-sensitive, embedding_service.py:452) file filter, so that
    item i describes embedding row i."""
    import uuid
    from PIL import Image
    items = []
    for name in sorted(f for f in os.listdir(folder_path) if f.endswith(IMAGE_EXT)):   # the rows of the ingest, exactly
        path = os.path.join(folder_path, name)
        with Image.open(path) as im:   # get_image_metadata, JSON_sample_DOC.py:12-26
            width, height = im.size
            metadata = {"size_bytes": os.path.getsize(path), "mime_type": Image.MIME.get(im.format, "unknown"),
                        "width": width, "height": height, "num_channels": len(im.getbands())}
        stem = os.path.splitext(name)[0]
        items.append({"id": str(uuid.uuid4()), "media_type": "image", "filepath": path, "tags": ["MainData"],
                      "metadata": metadata, "video": path_video, "frameid": name,
                      "text_detections": {"detections": []}, "object_detections": {"detections": []},
                      "frameidx": int(stem) if stem.isdigit() else None})
    os.makedirs(os.path.dirname(output_json_path) or ".", exist_ok=True)
    with open(output_json_path, "w", encoding="utf-8") as f:
        json.dump(items, f, ensure_ascii=False, indent=4)
    return output_json_path


def ingest_video(embedding_service, video_path, video_name=None, model_name="original", threshold=30.0,
                 metadata_builder=frame_metadata_items):
    """The processing half of `api_upload_video` (app.py:636-659) without Flask: scene-cut key-frames
    (`extract_frames_from_video(path, frame_dir, threshold=30.0)`, :638) -> embeddings (`:644`) -> metadata (`:647`) ->
    `video_mapping.json` entry with the reference's five fields (`:650-659`).  Returns the mapping entry.
    `metadata_builder(frame_dir, metadata_path, video_path)` is the hook for the reference's OCR / YOLO pass."""
    from . import scenecut
    ps = embedding_service.path_service
    video_name = video_name or os.path.splitext(os.path.basename(video_path))[0]
    frame_dir = os.path.join(ps.base_dir, video_name)
    os.makedirs(frame_dir, exist_ok=True)
    scenecut.extract_frames_from_video(video_path, frame_dir, threshold=threshold,
                                       device=str(embedding_service.device))
    if not any(f.endswith(IMAGE_EXT) for f in os.listdir(frame_dir)):
        raise RuntimeError(f"no scene cut found in {video_path} at threshold {threshold}: no key-frame to index")
    embeddings_file = embedding_service.extract_and_save_embeddings_from_folder(frame_dir, model_name, video_name)
    metadata_file = metadata_builder(frame_dir, os.path.join(ps.metadata_dir, f"{video_name}_metadata.json"),
                                     video_path)
    with open(metadata_file, "r", encoding="utf-8") as f:
        n_items = len(json.load(f))
    n_rows = np.load(embeddings_file, mmap_mode="r").shape[0]
    if n_items != n_rows:  # row i <-> item i is the identity every search result rests on
        raise RuntimeError(f"{video_name}: {n_rows} embedding rows but {n_items} metadata items")
    ps.add_video_mapping(video_name, metadata_file, embeddings_file, video_path, frame_dir)
    ps.video_data_mapping[video_name]["embedding_model"] = model_name
    ps.save_video_data_mapping()
    return ps.video_data_mapping[video_name]


def initialize_services(base_dir, metadata_dir, embedding_dir, finetuned_model_path=None, **embedding_kwargs):
    """services/__init__.py:11-64 for the services on this path (no visualisation service)."""
    path_service = PathService(base_dir, metadata_dir, embedding_dir)
    cache_service = CacheService()
    data_service = DataService(path_service, cache_service)
    embedding_service = EmbeddingService(cache_service, path_service, data_service, "cuda",
                                         checkpoint_path=finetuned_model_path, **embedding_kwargs)
    search_service = SearchService(embedding_service, data_service, path_service, cache_service)
    return {'path_service': path_service, 'cache_service': cache_service, 'data_service': data_service,
            'embedding_service': embedding_service, 'search_service': search_service, 'device': "cuda",
            'services_loaded': True}
