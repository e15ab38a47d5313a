"""Golden metadata items produced by the reference ITSELF: `get_image_metadata` and `process_image` are cut out of
/root/reference/Backend/JSON_sample_DOC.py (:12-87) with `ast` and run unmodified on the reference's own key-frames
(tests/golden/ref_frames/*.jpg) -- the module itself needs easyocr and ultralytics at import time.  The OCR reader and the
YOLO model are stand-ins that detect nothing (both are outside the hot path), so the items are what
`services.frame_metadata_items` must write: same keys, same values, `id` apart.  Authoring container only:
    python tests/golden/make_ref_metadata_golden.py
"""
import ast
import json
import os
import textwrap
import uuid

from PIL import Image, UnidentifiedImageError

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/Backend/JSON_sample_DOC.py"


def reference_functions():
    src = open(REF, encoding="utf-8").read()
    tree = ast.parse(src)
    ns = {"os": os, "Image": Image, "UnidentifiedImageError": UnidentifiedImageError, "uuid": uuid,
          "print": lambda *a, **k: None}
    for name in ("get_image_metadata", "process_image"):
        fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == name)
        exec(compile(textwrap.dedent(ast.get_source_segment(src, fn)), REF, "exec"), ns)
    return ns["process_image"]


class NoText:
    def readtext(self, path, detail=1):
        return []


class NoObjects:
    names = {}

    def __call__(self, path):
        return []


def main():
    process_image = reference_functions()
    frames = os.path.join(HERE, "ref_frames")
    items = []
    for name in sorted(f for f in os.listdir(frames) if f.endswith((".jpg", ".jpeg", ".png"))):
        item = process_image(os.path.join(frames, name), NoText(), "/videos/Violence.mp4", NoObjects())
        item["id"] = "<uuid4>"
        item["filepath"] = os.path.basename(item["filepath"])
        items.append(item)
    with open(os.path.join(HERE, "ref_metadata_golden.json"), "w", encoding="utf-8") as f:
        json.dump({"source": REF + ":12-87 (get_image_metadata, process_image, executed unmodified)", "items": items},
                  f, ensure_ascii=False, indent=1)
    print(len(items), "items;", items[0])


if __name__ == "__main__":
    main()
