class SentenceTransformer:  # names only
    pass
