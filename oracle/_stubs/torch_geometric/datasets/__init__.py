class MovieLens1M:  # names only
    pass
