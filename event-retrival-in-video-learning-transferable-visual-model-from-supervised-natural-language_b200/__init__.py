"""B200-native retrieval hot path: CLIP ViT encode_image + fused cosine scan/top-k (sm_100a CUDA behind a C ABI).

Drop-in mirror of the reference's Python surface for this path:
  clip            -- `clip.load / model.encode_image / model.visual.output_dim ...`  (openai/CLIP API)
  embedding       -- legacy `extract_and_save_embeddings_from_folder`                (Backend/embedding.py:9)
  services        -- `EmbeddingService`, `SearchService`, `initialize_services` and the path/cache/data services
  query_strategies-- the seven `query_by_*` strategies                                (Backend/query_strategies.py)
  metrics         -- R@K / MRR of the reference's compare_models harness through the scan  (compare_models.py:999-1090)
  app             -- `api_search(body, services)`: the /api/search endpoint without Flask (Backend/app.py:381-613)
  search          -- FrameStore: HBM-resident fp16 store + fused top-k (single GPU and row-sharded)
  vit             -- VisionEncoder: the hand-written tower behind `encode_image`
  text            -- TextEncoder: the text tower behind `encode_text`; tokenizer: `clip.tokenize` (BPE restated)
The CUDA library (lib/libevr.so, built by __graft_entry__.build()) is required; there is no CPU fallback.
"""
from . import _lib  # noqa: F401  (fails loudly if libevr.so is missing)
from . import search, text, vit  # noqa: F401
from . import app, clip, embedding, metrics, preprocess, query_strategies, scenecut, services, store  # noqa: F401

__all__ = ["search", "vit", "text", "clip", "embedding", "preprocess", "query_strategies", "scenecut", "services",
           "store", "app", "metrics", "_lib"]
