"""Type-only stand-in for anndata (the reference uses AnnData as a duck type)."""


class AnnData:
    pass
