"""`unidecode` for the strings this path sees (Latin script, Vietnamese diacritics).

The reference strips accents with the third-party `unidecode` package before every keyword / object substring test
(Backend/query_strategies.py:5, 206, 302, 383; Backend/services/search_service.py:7, 42).  The package is not installed
in this image.  When it is importable it is used; otherwise the stand-in below gives the same answer on Latin text:
canonical decomposition, combining marks dropped, and the handful of Latin letters that do not decompose mapped by
hand (d with stroke, o-slash, ...).  Characters outside Latin script pass through unchanged (unidecode would
transliterate them); OCR labels and YOLO class names of the reference's metadata are Vietnamese / English only.
"""
import unicodedata

try:  # the real thing, if a deployment has it
    from unidecode import unidecode as _real_unidecode
except ImportError:  # this image
    _real_unidecode = None

_SINGLE = {"đ": "d", "Đ": "D", "ø": "o", "Ø": "O", "ł": "l", "Ł": "L", "ß": "ss", "æ": "ae", "Æ": "AE", "œ": "oe",
           "Œ": "OE", "ı": "i", "ð": "d", "Ð": "D", "þ": "th", "Þ": "Th"}


def unaccent(text):
    out = []
    for ch in unicodedata.normalize("NFD", text):
        if unicodedata.combining(ch):
            continue
        out.append(_SINGLE.get(ch, ch))
    return "".join(out)


def unidecode(text):
    return _real_unidecode(text) if _real_unidecode is not None else unaccent(text)
