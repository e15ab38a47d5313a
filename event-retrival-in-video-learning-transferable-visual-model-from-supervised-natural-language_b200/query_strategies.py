"""The seven query strategies over the service callbacks.

Mirror of Backend/query_strategies.py with the same names, positional signatures, keyword defaults and return values:
    query_by_text_clip :36                     query_by_text_with_adaptive_threshold :121
    query_by_keyword :188                      query_by_text_and_keyword :253
    query_by_object :363                       query_by_text_and_object :466
    query_by_text_object_and_keyword :601
The CLIP-using ones call `search_top_frames(text, top_k*3, video)` (the fused scan) and `extract_query_confidence`
(served from that scan's scores) through the callbacks `api_search` hands in, join each candidate with its metadata
record, apply the thresholds, build the event with `format_event_for_frontend`, sort, `[:top_k]`; `[]` on any error.
The OCR-keyword / object tests are substring matches on accent-stripped lower-case labels, as in the reference.

What differs on the host side (results are the same; tests/test_query_surface.py replays the reference's own
functions against these):
  * the metadata JSON is parsed once per (path, mtime) and indexed by `frameidx` and `frameid`; the reference re-reads
    2 MB+ per request and walks it with `next(...)` per candidate (:59-63, :95);
  * the Vietnamese preprocessing / translation step (word_processing.VietnameseTextProcessor: pyvi, underthesea, a
    translation RPC -- none available offline) is an injectable `text_processor` (default: identity);
  * the fallback metadata path is `METADATA_DIR/output_samples.json` with METADATA_DIR settable instead of the
    hard-coded E:\\ path (:9-10).
"""
import json
import os
import threading
from pathlib import Path

from ._unaccent import unidecode

METADATA_DIR = os.environ.get("EVR_METADATA_DIR", os.path.join(os.getcwd(), "metadata"))
text_processor = None  # callable str -> str, e.g. VietnameseTextProcessor().preprocess_and_translate

CAPTION_CONFIDENCE = 0.65   # :429  a caption hit counts as this much object confidence
TAG_CONFIDENCE = 0.75       # :440
OCR_AS_OBJECT_SCALE = 0.7   # :454  an object name read by OCR is worth 70 % of the OCR confidence
OBJECT_THRESHOLD_CAP = 0.65  # :461 query_by_object never asks for more than this

_meta_cache = {}
_meta_lock = threading.Lock()


def get_default_metadata_path():
    return os.path.join(METADATA_DIR, "output_samples.json")


def get_video_metadata_path(video_name=None, video_data_mapping=None):
    if video_name and video_data_mapping and video_name in video_data_mapping:
        return os.path.normpath(video_data_mapping[video_name]["metadata_file"])
    return get_default_metadata_path()


class _Records:
    """One metadata file: the records in file order plus first-match indexes (what `next(...)` finds)."""

    def __init__(self, data):
        self.data = data
        self.by_frameidx, self.by_frameid = {}, {}
        for item in data:
            self.by_frameidx.setdefault(item.get('frameidx'), item)
            self.by_frameid.setdefault(item.get('frameid'), item)


def _records(video_name, video_data_mapping):
    json_path = get_video_metadata_path(video_name, video_data_mapping)
    mtime = os.path.getmtime(json_path)
    with _meta_lock:
        hit = _meta_cache.get(json_path)
        if hit is not None and hit[0] == mtime:
            return hit[1]
    with open(json_path, "r", encoding="utf-8") as f:
        rec = _Records(json.load(f))
    with _meta_lock:
        if len(_meta_cache) >= 64:
            _meta_cache.clear()
        _meta_cache[json_path] = (mtime, rec)
    return rec


def _metadata_by_frameidx(json_path):
    """{frameidx: first record}; kept for callers that hold a path (services.SearchService)."""
    with open(json_path, "r", encoding="utf-8") as f:
        return _Records(json.load(f)).by_frameidx


def _process(query):
    return text_processor(query) if text_processor is not None else query


def _plain(x):
    """numpy scalars -> python floats (the reference converts np.float32 by hand, e.g. :296-298)."""
    return float(x) if type(x).__module__ == "numpy" else x


# ---- metadata matchers ------------------------------------------------------------------------------------------
def _ocr_confidence(record, keyword):
    """Highest confidence among the OCR detections whose accent-stripped label contains the accent-stripped keyword;
    0 when none matches (a match at confidence 0 counts as none, as `conf > max_conf` does at :211, :318)."""
    needle = unidecode(keyword.lower())
    best = 0
    for det in record.get("text_detections", {}).get("detections", []):
        if needle in unidecode(det.get("label", "").lower()):
            conf = det.get("confidence", 0)
            if conf > best:
                best = conf
    return best


def _object_confidence(record, name, ocr_too):
    """(confidence, label) of the best evidence that `name` is in the frame: YOLO detections, then the caption (0.65),
    the tags (0.75) and -- where the reference looks there (`ocr_too`) -- OCR text at 70 % (:411-458, :521-557,
    :683-728).  (0, "") when there is none."""
    low = name.lower()
    flat = unidecode(low)

    def hit(text):
        return low in text or flat in unidecode(text)

    best, label = 0, ""
    for obj in record.get("object_detections", {}).get("detections", []):
        obj_label = obj.get("label", "").lower()
        if hit(obj_label):
            conf = obj.get("confidence", 0)
            if conf > best:
                best, label = conf, obj_label
    meta = record.get("metadata", {})
    caption = meta.get("caption", "").lower() if meta else ""
    if caption and hit(caption) and CAPTION_CONFIDENCE > best:
        best, label = CAPTION_CONFIDENCE, name
    for tag in record.get("tags", []):
        if hit(str(tag).lower()) and TAG_CONFIDENCE > best:
            best, label = TAG_CONFIDENCE, tag
    if ocr_too:
        for det in record.get("text_detections", {}).get("detections", []):
            det_label = det.get("label", "").lower()
            if hit(det_label):
                conf = det.get("confidence", 0) * OCR_AS_OBJECT_SCALE
                if conf > best:
                    best, label = conf, det_label
    return best, label


def _clip_candidates(query, top_k, search_top_frames, video_name, video_data_mapping):
    processed_text = _process(query)
    query_frames = search_top_frames(processed_text, top_k * 3, video_name)
    return processed_text, query_frames, _records(video_name, video_data_mapping)


def _scored(frame_name, rec, processed_text, video_name, extract_query_confidence):
    """(record, cosine) of one candidate in the combined strategies; (None, None) when it has no metadata record.
    The store searched is the request's video, else the one the record names (:283-289)."""
    record = rec.by_frameidx.get(int(Path(frame_name).stem))
    if not record:
        return None, None
    video_path = record.get('video', '')
    effective_video = video_name or (Path(video_path).stem if video_path else None)
    return record, _plain(extract_query_confidence(frame_name, processed_text, effective_video))


# ---- strategies -------------------------------------------------------------------------------------------------
def query_by_text_clip(query, top_k, search_top_frames, extract_query_confidence, format_event_for_frontend,
                       video_name=None, video_data_mapping=None):
    try:
        processed_text, query_frames, rec = _clip_candidates(query, top_k, search_top_frames, video_name,
                                                             video_data_mapping)
        results, seen = [], set()
        for frame_name in query_frames:
            if frame_name in seen:
                continue
            seen.add(frame_name)
            try:
                record = rec.by_frameidx.get(int(Path(frame_name).stem))
            except Exception:
                continue
            if not record:
                continue
            try:
                confidence = extract_query_confidence(frame_name, processed_text, video_name)
                record = dict(record, clip_similarity=confidence)
                event = format_event_for_frontend(record)
                event["clip_similarity"] = confidence
                results.append(event)
            except Exception as e:
                print(f"Error processing frame {frame_name}: {e}")
        results.sort(key=lambda x: x.get("clip_similarity", 0), reverse=True)
        return results[:top_k]
    except Exception as e:
        print(f"Error in text clip query: {e}")
        return []


def query_by_text_with_adaptive_threshold(query, adaptive_threshold, top_k, search_top_frames,
                                          extract_query_confidence, format_event_for_frontend, video_name=None,
                                          video_data_mapping=None):
    try:
        processed_text, query_frames, rec = _clip_candidates(query, top_k, search_top_frames, video_name,
                                                             video_data_mapping)
        results = []
        for frame_name in query_frames:
            try:
                record = rec.by_frameidx.get(int(Path(frame_name).stem))
                if not record:
                    continue
                confidence = extract_query_confidence(frame_name, processed_text, video_name)
                if confidence >= adaptive_threshold:
                    event = format_event_for_frontend(dict(record, clip_similarity=confidence))
                    event["clip_similarity"] = confidence
                    results.append(event)
            except Exception as e:
                print(f"Error processing frame {frame_name}: {e}")
        results.sort(key=lambda x: x.get("clip_similarity", 0), reverse=True)
        return results[:top_k]
    except Exception as e:
        print(f"Error in query_by_text_with_adaptive_threshold: {e}")
        return []


def query_by_keyword(query, adaptive_threshold, top_k, search_frames_by_keyword, format_event_for_frontend,
                     video_name=None, video_data_mapping=None):
    try:
        keyword_frame_ids = search_frames_by_keyword(query, top_k * 3, video_name)
        rec = _records(video_name, video_data_mapping)
        results = []
        for frame_id in keyword_frame_ids:
            record = rec.by_frameid.get(frame_id)
            if not record:
                continue
            conf = _ocr_confidence(record, query)
            if conf and conf >= adaptive_threshold:
                results.append(format_event_for_frontend(dict(record, text_confidence=conf, clip_similarity=0.0)))
        results.sort(key=lambda x: x.get("confidence", 0), reverse=True)
        return results[:top_k]
    except Exception as e:
        print(f"Error in keyword query: {e}")
        return []


def query_by_text_and_keyword(query, adaptive_threshold, top_k, search_top_frames, extract_query_confidence,
                              search_frames_by_keyword, format_event_for_frontend, keyword=None, text_confidence=None,
                              video_name=None, video_data_mapping=None):
    keyword = keyword if keyword else query
    keyword_threshold = text_confidence if text_confidence is not None else adaptive_threshold
    try:
        processed_text, query_frames, rec = _clip_candidates(query, top_k, search_top_frames, video_name,
                                                             video_data_mapping)
        keyword_frame_ids = search_frames_by_keyword(keyword, top_k * 3, video_name)
        results = []
        for frame_name in query_frames:
            record, clip_conf = _scored(frame_name, rec, processed_text, video_name, extract_query_confidence)
            if record is None or not clip_conf >= adaptive_threshold:
                continue
            if record.get('frameid') not in keyword_frame_ids:
                continue
            ocr_conf = _ocr_confidence(record, keyword)
            if not (ocr_conf and ocr_conf >= keyword_threshold):
                continue
            fields = dict(record, clip_similarity=clip_conf, text_confidence=_plain(ocr_conf))
            fields.setdefault('object_confidence', 0.0)
            event = format_event_for_frontend(fields)
            event.update(confidence=max(clip_conf, ocr_conf), clip_similarity=clip_conf, text_confidence=ocr_conf,
                         detection_type="text+clip")
            results.append(event)
        results.sort(key=lambda x: x["confidence"], reverse=True)
        return results[:top_k]
    except Exception as e:
        print(f"Error in text and keyword query: {e}")
        return []


def query_by_object(query, adaptive_threshold, top_k, format_event_for_frontend, video_name=None,
                    video_data_mapping=None):
    try:
        rec = _records(video_name, video_data_mapping)
        threshold = min(adaptive_threshold, OBJECT_THRESHOLD_CAP)
        results = []
        for record in rec.data:
            conf, label = _object_confidence(record, query, ocr_too=True)
            if conf and conf >= threshold:
                fields = dict(record, object_confidence=_plain(conf), detection_type='object')
                if not fields.get('object_label'):
                    fields['object_label'] = label
                results.append(format_event_for_frontend(fields))
        results.sort(key=lambda x: x["confidence"], reverse=True)
        return results[:top_k]
    except Exception as e:
        print(f"Error in object search: {e}")
        return []


def query_by_text_and_object(query, adaptive_threshold, top_k, search_top_frames, extract_query_confidence,
                             format_event_for_frontend, object_keyword=None, object_confidence=None,
                             video_name=None, video_data_mapping=None):
    object_name = object_keyword if object_keyword else query
    object_threshold = object_confidence if object_confidence is not None else adaptive_threshold
    try:
        processed_text, query_frames, rec = _clip_candidates(query, top_k, search_top_frames, video_name,
                                                             video_data_mapping)
        results = []
        for frame_name in query_frames:
            record, clip_conf = _scored(frame_name, rec, processed_text, video_name, extract_query_confidence)
            if record is None or not clip_conf >= adaptive_threshold:
                continue
            obj_conf, label = _object_confidence(record, object_name, ocr_too=False)   # no OCR fallback here (:521-557)
            if not (obj_conf and obj_conf >= object_threshold):
                continue
            fields = dict(record, clip_similarity=clip_conf, object_confidence=_plain(obj_conf), object_label=label)
            fields.setdefault('text_confidence', 0.0)
            event = format_event_for_frontend(fields)
            event.update(confidence=max(clip_conf, obj_conf), clip_similarity=clip_conf, object_confidence=obj_conf,
                         detection_type="object+clip")
            results.append(event)
        results.sort(key=lambda x: x["confidence"], reverse=True)
        return results[:top_k]
    except Exception as e:
        print(f"Error in text and object query: {e}")
        return []


def query_by_text_object_and_keyword(query, adaptive_threshold, top_k, search_top_frames, extract_query_confidence,
                                     search_frames_by_keyword, format_event_for_frontend, keyword=None,
                                     text_confidence=None, object_keyword=None, object_confidence=None,
                                     video_name=None, video_data_mapping=None):
    keyword = keyword if keyword else query
    object_name = object_keyword if object_keyword else query
    keyword_threshold = text_confidence if text_confidence is not None else adaptive_threshold
    object_threshold = object_confidence if object_confidence is not None else adaptive_threshold
    try:
        processed_text, query_frames, rec = _clip_candidates(query, top_k, search_top_frames, video_name,
                                                             video_data_mapping)
        keyword_frame_ids = search_frames_by_keyword(keyword, top_k * 3, video_name)
        results = []
        for frame_name in query_frames:
            record, clip_conf = _scored(frame_name, rec, processed_text, video_name, extract_query_confidence)
            if record is None or not clip_conf >= adaptive_threshold:
                continue
            if record.get('frameid') not in keyword_frame_ids:
                continue
            ocr_conf = _ocr_confidence(record, keyword)
            if not (ocr_conf and ocr_conf >= keyword_threshold):
                continue
            obj_conf, label = _object_confidence(record, object_name, ocr_too=True)
            if not (obj_conf and obj_conf >= object_threshold):
                continue
            event = format_event_for_frontend(dict(record, clip_similarity=clip_conf, text_confidence=ocr_conf,
                                                   object_confidence=obj_conf, object_label=label))
            event.update(confidence=max(clip_conf, ocr_conf, obj_conf), clip_similarity=clip_conf,
                         text_confidence=ocr_conf, object_confidence=obj_conf, detection_type="text+object+clip")
            results.append(event)
        results.sort(key=lambda x: x["confidence"], reverse=True)
        return results[:top_k]
    except Exception as e:
        print(f"Error in text, object and keyword query: {e}")
        return []


def search_by_image(embedding_service, image_url, adaptive_threshold, top_k, format_event_for_frontend, video_name=None,
                    video_data_mapping=None):
    """Round-1 free-function form of the image query; `services.SearchService.search_by_image` is the drop-in
    (reference signature).  Kept as a thin forwarder for callers that hold no SearchService."""
    from .services import SearchService
    svc = SearchService(embedding_service, embedding_service.data_service, embedding_service.path_service,
                        embedding_service.cache_service)
    return svc.search_by_image(image_url, adaptive_threshold, top_k, video_name,
                               format_event_for_frontend=format_event_for_frontend)
