"""Text-query strategies over the two service callbacks.

Mirror of Backend/query_strategies.py `query_by_text_clip` (:36-119) and `query_by_text_with_adaptive_threshold`
(:121-186): same positional signatures, same steps -- `search_top_frames(text, top_k*3, video)`, join each candidate
with its metadata record by `frameidx == int(stem)`, `extract_query_confidence`, [threshold,] `format_event_for_frontend`,
sort by `clip_similarity`, `[:top_k]`; `[]` on any error.  The callbacks are `EmbeddingService.search_top_frames` and
`.extract_query_confidence`, i.e. the fused scan and `evr_rescore`.

Host-side differences: the metadata JSON is parsed once per (path, mtime) and indexed by frameidx (the reference
re-reads 2 MB+ and walks it with `next(...)` per candidate); the Vietnamese preprocessing / translation step
(word_processing.VietnameseTextProcessor: pyvi, underthesea, a translation RPC -- none available offline) is an
injectable `text_processor` (default: identity); the metadata fallback path is `METADATA_DIR/output_samples.json`
with METADATA_DIR settable instead of the hard-coded E:\\ path (:9-10).
"""
import json
import os
import threading
from pathlib import Path

METADATA_DIR = os.environ.get("EVR_METADATA_DIR", os.path.join(os.getcwd(), "metadata"))
text_processor = None  # callable str -> str, e.g. VietnameseTextProcessor().preprocess_and_translate

_meta_cache = {}
_meta_lock = threading.Lock()


def get_default_metadata_path():
    return os.path.join(METADATA_DIR, "output_samples.json")


def get_video_metadata_path(video_name=None, video_data_mapping=None):
    if video_name and video_data_mapping and video_name in video_data_mapping:
        return os.path.normpath(video_data_mapping[video_name]["metadata_file"])
    return get_default_metadata_path()


def _metadata_by_frameidx(json_path):
    """{frameidx: first record with it} -- what `next(item for item in data if item.get('frameidx') == idx)` finds."""
    mtime = os.path.getmtime(json_path)
    with _meta_lock:
        hit = _meta_cache.get(json_path)
        if hit is not None and hit[0] == mtime:
            return hit[1]
    with open(json_path, "r", encoding="utf-8") as f:
        data = json.load(f)
    table = {}
    for item in data:
        table.setdefault(item.get('frameidx'), item)
    with _meta_lock:
        _meta_cache[json_path] = (mtime, table)
    return table


def _process(query):
    return text_processor(query) if text_processor is not None else query


def _events(query_frames, table, processed_text, video_name, extract_query_confidence, format_event_for_frontend,
            threshold=None):
    results, seen = [], set()
    for frame_name in query_frames:
        if frame_name in seen:
            continue
        seen.add(frame_name)
        try:
            frame_data = table.get(int(Path(frame_name).stem))
            if not frame_data:
                continue
            confidence = extract_query_confidence(frame_name, processed_text, video_name)
            if threshold is not None and not confidence >= threshold:
                continue
            frame_data_copy = frame_data.copy()
            frame_data_copy['clip_similarity'] = confidence
            event = format_event_for_frontend(frame_data_copy)
            event["clip_similarity"] = confidence
            results.append(event)
        except Exception as e:
            print(f"Error processing frame {frame_name}: {e}")
    results.sort(key=lambda x: x.get("clip_similarity", 0), reverse=True)
    return results


def query_by_text_clip(query, top_k, search_top_frames, extract_query_confidence, format_event_for_frontend,
                       video_name=None, video_data_mapping=None):
    try:
        processed_text = _process(query)
        query_frames = search_top_frames(processed_text, top_k * 3, video_name)
        table = _metadata_by_frameidx(get_video_metadata_path(video_name, video_data_mapping))
        return _events(query_frames, table, processed_text, video_name, extract_query_confidence,
                       format_event_for_frontend)[:top_k]
    except Exception as e:
        print(f"Error in text clip query: {e}")
        return []


def query_by_text_with_adaptive_threshold(query, adaptive_threshold, top_k, search_top_frames,
                                          extract_query_confidence, format_event_for_frontend, video_name=None,
                                          video_data_mapping=None):
    try:
        processed_text = _process(query)
        query_frames = search_top_frames(processed_text, top_k * 3, video_name)
        table = _metadata_by_frameidx(get_video_metadata_path(video_name, video_data_mapping))
        return _events(query_frames, table, processed_text, video_name, extract_query_confidence,
                       format_event_for_frontend, threshold=adaptive_threshold)[:top_k]
    except Exception as e:
        print(f"Error in query_by_text_with_adaptive_threshold: {e}")
        return []


def search_by_image(embedding_service, image_url, adaptive_threshold, top_k, format_event_for_frontend, video_name=None,
                    video_data_mapping=None):
    """Image-query path (Backend/services/search_service.py:611-705): encode the query image,
    scan with `top_k * 3`, keep candidates whose cosine passes `adaptive_threshold`, best `top_k` events.

    `image_url`: file path, `data:image/...;base64,` URL, or a PIL image (http(s) URLs need the network and are
    refused).  Difference from the reference: a candidate's similarity is the scan's own score -- the same
    dot product of its store row with the query -- instead of re-embedding the frame with
    `extract_image_embedding(<basename>)`, which fails outside the frames directory and silently drops the candidate
    (search_service.py:679)."""
    try:
        import base64
        from io import BytesIO

        import numpy as np
        import torch
        from PIL import Image
        if hasattr(image_url, "convert"):
            img = image_url.convert('RGB')
        elif image_url.startswith(('http://', 'https://')):
            raise RuntimeError("remote image URLs need network access")
        elif image_url.startswith('data:image'):
            img = Image.open(BytesIO(base64.b64decode(image_url.split(',')[1]))).convert('RGB')
        else:
            img = Image.open(image_url).convert('RGB')
        svc = embedding_service
        image = svc.preprocess(img).unsqueeze(0).to(svc.device)
        with torch.no_grad():
            if svc.get_active_model_name() == "finetuned" and svc.finetuned_model is not None:
                feats = svc.finetuned_model(image)
            else:
                feats = svc.original_model.encode_image(image)
            feats = feats.float()
            feats = (feats / feats.norm(dim=-1, keepdim=True)).cpu().numpy()
        table = _metadata_by_frameidx(get_video_metadata_path(video_name, video_data_mapping))
        results = []
        for frame_name, similarity in svc.search_with_scores(feats, top_k * 3, video_name):
            try:
                frame_data = table.get(int(Path(frame_name).stem))
                if frame_data and similarity >= adaptive_threshold:
                    rec = frame_data.copy()
                    rec['clip_similarity'] = float(similarity)
                    event = format_event_for_frontend(rec)
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
