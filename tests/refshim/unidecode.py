"""Stand-in for the absent third-party `unidecode` package so that the reference's modules import unmodified.
Same rule as the product's `_unaccent.unaccent` (NFD, drop combining marks, a few stroke letters by hand), restated
here so the shim does not import the product."""
import unicodedata

_SINGLE = {"đ": "d", "Đ": "D", "ø": "o", "Ø": "O", "ł": "l", "Ł": "L", "ß": "ss", "æ": "ae", "Æ": "AE", "œ": "oe",
           "Œ": "OE", "ı": "i", "ð": "d", "Ð": "D", "þ": "th", "Þ": "Th"}


def unidecode(text):
    return "".join(_SINGLE.get(c, c) for c in unicodedata.normalize("NFD", text) if not unicodedata.combining(c))
