from oracle.pyg_semantics import add_self_loops, remove_self_loops, segment_softmax


def softmax(src, index, ptr=None, num_nodes=None, dim=0):
    n = int(index.max()) + 1 if num_nodes is None else num_nodes
    return segment_softmax(src, index, n)
