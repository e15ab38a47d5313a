"""Stand-in for Backend/word_processing.py, whose imports (pyvi, underthesea, langdetect, deep_translator + a
translation RPC) are unavailable offline: the preprocessing step becomes the identity, which is also the product's
default `query_strategies.text_processor`."""


class VietnameseTextProcessor:
    def preprocess_and_translate(self, text):
        return text
