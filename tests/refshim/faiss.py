"""Empty stand-in: the reference imports faiss (embedding_service.py:10, app.py:21) and never calls it."""
