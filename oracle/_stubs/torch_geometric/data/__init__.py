class HeteroData:  # names only
    pass


class InMemoryDataset:
    pass


def download_google_url(*a, **k):
    raise RuntimeError("stub")


def extract_zip(*a, **k):
    raise RuntimeError("stub")


def download_url(*a, **k):
    raise RuntimeError("stub")
