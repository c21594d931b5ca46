"""Batches handed to user-defined functions (python/dgl/udf.py:8-181 in the reference):
EdgeBatch for message / apply-edge UDFs, NodeBatch for reduce / apply-node UDFs."""


class EdgeBatch(object):
    def __init__(self, g, edges, src_data, edge_data, dst_data):
        self._g = g
        self._edges = edges
        self._src_data = src_data
        self._edge_data = edge_data
        self._dst_data = dst_data

    @property
    def src(self):
        return self._src_data

    @property
    def dst(self):
        return self._dst_data

    @property
    def data(self):
        return self._edge_data

    def edges(self):
        """(u, v, eid) of the batch."""
        return self._edges

    def batch_size(self):
        return int(self._edges[0].shape[0])

    def __len__(self):
        return self.batch_size()


class NodeBatch(object):
    def __init__(self, g, nodes, data, msgs=None):
        self._g = g
        self._nodes = nodes
        self._data = data
        self._msgs = msgs

    @property
    def data(self):
        return self._data

    @property
    def mailbox(self):
        return self._msgs

    def nodes(self):
        return self._nodes

    def batch_size(self):
        if self._nodes is None:
            return self._g.number_of_nodes()
        return int(self._nodes.shape[0])

    def __len__(self):
        return self.batch_size()
