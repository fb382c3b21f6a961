from oracle.pyg_semantics import glorot_


def glorot(t):
    if t is not None:
        glorot_(t.data if hasattr(t, "data") else t)


def zeros(t):
    if t is not None:
        t.data.fill_(0)
