"""The search endpoint without Flask: `api_search(body, services) -> (payload, status)`.

Mirror of `api_search()` in Backend/app.py:381-613 -- the function that produces the ranked events the user sees:
    body parsing with the reference's defaults                      :386-423
    `set_active_model` when the request names another model         :410-412
    "video-<n>" -> n-th key of video_mapping.json                   :414-418
    dispatch on search_type / search_method                         :434-555
    missing confidence fields filled with 0.0                       :558-564
    per-video filter on the event's videoId                         :567-588
    final sort: clip_similarity for image / text_clip / text_adaptive / enableClipSimilarity, else confidence  :591-598
    {"events": results[:top_k]}                                      :609   ({"error": str(e)}, 500 on an exception :611-613)
and of the forwarding functions the dispatch calls (app.py:151-219).  The HTTP layer (Flask routing, `request.json`,
`jsonify`) is out of scope: a server hands the decoded JSON body in and serialises the returned dict.

`services` is the dict `initialize_services(...)` returns (services/__init__.py:57-65).  Every CLIP call underneath
(`search_top_frames`, `extract_query_confidence`, `search_by_image`) is the fused scan / the sm_100a tower.
"""
import time
from pathlib import Path

from . import query_strategies as qs

CLIP_SORTED_METHODS = ("text_clip", "text_adaptive")  # :591


class SearchApp:
    """The module-level state of app.py this path reads: the service objects and the forwarding functions."""

    def __init__(self, services):
        self.path_service = services['path_service']
        self.cache_service = services['cache_service']
        self.data_service = services['data_service']
        self.embedding_service = services['embedding_service']
        self.search_service = services['search_service']

    @property
    def video_data_mapping(self):
        return self.path_service.video_data_mapping

    # ---- forwarding functions (app.py:151-219) ------------------------------------------------------
    def format_event_for_frontend(self, frame_data):
        return self.data_service.format_event_for_frontend(frame_data)

    def search_frames_by_keyword(self, keyword, top_k, video_name=None):
        return self.search_service.search_frames_by_keyword(keyword, top_k, video_name)

    def extract_query_confidence(self, frame_path, query, video_name=None):
        return self.embedding_service.extract_query_confidence(frame_path, query, video_name)

    def search_top_frames(self, query, top_k, video_name=None):
        return self.embedding_service.search_top_frames(query, top_k, video_name)

    def search_top_frames_by_image(self, image_features, top_k, video_name=None):
        return self.embedding_service.search_top_frames_by_image(image_features, top_k, video_name)

    def search_by_image(self, image_url, adaptive_threshold, top_k, video_name=None):
        return self.search_service.search_by_image(image_url, adaptive_threshold, top_k, video_name,
                                                   self.embedding_service.preprocess)

    # ---- the endpoint --------------------------------------------------------------------------------
    def api_search(self, data):
        try:
            start_time = time.time()
            search_type = data.get("search_type", "text")
            query = data.get("query", "")
            image_url = data.get("image_url")
            top_k = data.get("top_k", 10)
            adaptive_threshold = float(data.get("adaptive_threshold", 0.5))
            text_confidence = float(data.get("text_confidence", adaptive_threshold))
            object_confidence = float(data.get("object_confidence", adaptive_threshold))
            search_method = data.get("search_method", "text")
            keyword = data.get("keyword", "")
            object_keyword = data.get("object", "")
            video_id = data.get("videoId")
            model_name = data.get("model", "original")
            enable_clip_similarity = data.get("enableClipSimilarity", False)

            if model_name != self.embedding_service.get_active_model_name():
                self.embedding_service.set_active_model(model_name)

            mapping = self.video_data_mapping
            video_name = None
            if video_id and video_id.startswith('video-'):
                video_num = int(video_id.split('-')[1])
                all_videos = list(mapping.keys())
                if all_videos and len(all_videos) >= video_num:
                    video_name = all_videos[video_num - 1]

            clip = (self.search_top_frames, self.extract_query_confidence)
            where = dict(video_name=video_name, video_data_mapping=mapping)
            results = []
            if search_type == "image" and image_url:
                results = self.search_by_image(image_url, adaptive_threshold, top_k, video_name)
            elif search_type == "text" and query:
                if search_method == "text_clip":
                    results = qs.query_by_text_clip(query, top_k, *clip, self.format_event_for_frontend, **where)
                elif search_method == "keyword_only":
                    results = qs.query_by_keyword(keyword if keyword else query, text_confidence, top_k,
                                                  self.search_frames_by_keyword, self.format_event_for_frontend,
                                                  **where)
                elif search_method == "text_keyword":
                    results = qs.query_by_text_and_keyword(
                        query, adaptive_threshold, top_k, *clip, self.search_frames_by_keyword,
                        self.format_event_for_frontend, keyword=keyword if keyword else query,
                        text_confidence=text_confidence, **where)
                elif search_method == "object_only":
                    results = qs.query_by_object(object_keyword if object_keyword else query, object_confidence, top_k,
                                                 self.format_event_for_frontend, **where)
                elif search_method == "text_object":
                    results = qs.query_by_text_and_object(
                        query, adaptive_threshold, top_k, *clip, self.format_event_for_frontend,
                        object_keyword=object_keyword if object_keyword else query,
                        object_confidence=object_confidence, **where)
                elif search_method == "text_object_keyword":
                    results = qs.query_by_text_object_and_keyword(
                        query, adaptive_threshold, top_k, *clip, self.search_frames_by_keyword,
                        self.format_event_for_frontend, keyword=keyword if keyword else query,
                        text_confidence=text_confidence, object_keyword=object_keyword if object_keyword else query,
                        object_confidence=object_confidence, **where)
                else:  # "text_adaptive" and every unknown method (:545-555)
                    results = qs.query_by_text_with_adaptive_threshold(query, adaptive_threshold, top_k, *clip,
                                                                       self.format_event_for_frontend, **where)

            for result in results:
                for field in ("text_confidence", "object_confidence", "clip_similarity"):
                    if field not in result:
                        result[field] = 0.0

            if video_name:
                video_path = mapping.get(video_name, {}).get("video_path")
                if video_path:
                    stem = Path(video_path).stem
                    wanted = (f"video-{video_name}", f"video-{stem}")
                    results = [r for r in results if r.get("videoId") in wanted or stem in r.get("videoId")]

            if search_type == "image" or search_method in CLIP_SORTED_METHODS or enable_clip_similarity:
                results = sorted(results, key=lambda x: x.get("clip_similarity", 0), reverse=True)
            else:
                results = sorted(results, key=lambda x: x.get("confidence", 0), reverse=True)
            self.last_search_seconds = time.time() - start_time
            return {"events": results[:top_k]}, 200
        except Exception as e:
            print(f"Error searching: {e}")
            return {"error": str(e)}, 500


def api_search(body, services):
    """One `/api/search` request: decoded JSON body -> (response dict, HTTP status)."""
    app = services.get('_search_app') if isinstance(services, dict) else None
    if app is None:
        app = SearchApp(services)
        if isinstance(services, dict):
            services['_search_app'] = app
    return app.api_search(body)
