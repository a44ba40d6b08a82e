"""`KNNGraph` / `KNNEdges`: the initial per-event kNN graph, built on the device.

Mirrors the reference's call contract
(/root/reference/src/graphnet/models/data_representation/graphs/graphs.py:21-73,
graph_definition.py:127-184, edges/edges.py:51-84): a `KNNGraph` is called with
one event's `[n, d]` numpy feature array and returns a `Data` whose `x` is
standardised by the detector and whose `edge_index` is
`knn_graph(x[:, columns], nb_nearest_neighbours)`.

Two additions for the GPU: `KNNGraph.collate(...)` builds the
graphs of a whole batch of events with ONE kernel launch after collation, which
is how the dataset-side CPU KD-tree is taken off the critical path; and
`KNNEdges` accepts a `Data` that already carries `batch`.
"""

from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import torch
from torch import Tensor

from . import ops
from .data import Batch, Data
from .detector import Detector


class KNNEdges:
    """edges.py:51-84."""

    def __init__(self, nb_nearest_neighbours: int, columns: List[int] = [0, 1, 2]):
        self._nb_nearest_neighbours = nb_nearest_neighbours
        self._columns = list(columns)

    def __call__(self, graph: Data) -> Data:
        return self._construct_edges(graph)

    def _construct_edges(self, graph: Data) -> Data:
        batch = getattr(graph, "batch", None)
        graph.edge_index = ops.knn_graph(
            graph.x[:, self._columns], self._nb_nearest_neighbours, batch)
        return graph


class KNNGraph:
    """graphs.py:21-73 (NodesAsPulses node definition, KNNEdges edge definition)."""

    def __init__(
        self,
        detector: Detector,
        node_definition: Any = None,
        input_feature_names: Optional[List[str]] = None,
        dtype: Optional[torch.dtype] = torch.float,
        perturbation_dict: Optional[Dict[str, float]] = None,
        seed: Optional[int] = None,
        nb_nearest_neighbours: int = 8,
        columns: List[int] = [0, 1, 2],
        distance_as_edge_feature: bool = False,
        device: str = "cuda",
        **kwargs: Any,
    ) -> None:
        if distance_as_edge_feature:
            # graphs.py:57-59 -> KNNDistanceEdges (undirected edges + a distance attribute):
            # DynEdge reads neither; not part of this path
            raise NotImplementedError("distance_as_edge_feature=True (KNNDistanceEdges) is not built")
        if node_definition is not None:
            raise NotImplementedError("only the default NodesAsPulses node definition is built")
        self._detector = detector
        self._input_feature_names = input_feature_names
        self._dtype = dtype
        self._perturbation_dict = perturbation_dict
        self.rng = np.random.default_rng(seed)
        self._edge_definition = KNNEdges(nb_nearest_neighbours, columns)
        self._device = torch.device(device)

    @property
    def nb_inputs(self) -> int:
        return len(self._input_feature_names) if self._input_feature_names else 0

    def _standardise(self, input_features: np.ndarray, names: List[str]) -> Tensor:
        feats = np.array(input_features, dtype=np.float64, copy=True)
        if self._perturbation_dict:  # data_representation.py:150-163
            for name, std in self._perturbation_dict.items():
                idx = names.index(name)
                feats[:, idx] = self.rng.normal(feats[:, idx], std)
        x = torch.tensor(feats, dtype=self._dtype)
        return self._detector(x, names)

    def __call__(self, input_features: np.ndarray, input_feature_names: List[str],
                 truth_dicts: Optional[List[Dict[str, Any]]] = None, **kwargs: Any) -> Data:
        x = self._standardise(input_features, input_feature_names).to(self._device)
        data = Data(x=x)
        data.n_pulses = torch.tensor(x.shape[0], dtype=torch.int32)  # data_representation.py:175
        if truth_dicts:
            for d in truth_dicts:
                for k, v in d.items():
                    data[k] = torch.tensor(v)
        data = self._edge_definition(data)
        for i, name in enumerate(input_feature_names):  # graph_definition.py:186-202
            data[name] = data.x[:, i].detach()
        return data

    def collate(self, events: Sequence[np.ndarray], input_feature_names: List[str]) -> Batch:
        """Standardise on the host, then ONE device kNN launch for all events."""
        xs = [self._standardise(e, input_feature_names) for e in events]
        keep = [x for x in xs if x.shape[0] > 1]  # collate_fn, data/dataloader.py:11-17
        counts = torch.tensor([x.shape[0] for x in keep], dtype=torch.int32)
        x = torch.cat(keep).to(self._device)
        batch = torch.repeat_interleave(torch.arange(len(keep)), counts.long()).to(self._device)
        out = Batch(x=x, batch=batch, n_pulses=counts.to(self._device))
        out.num_graphs = len(keep)
        out.edge_index = ops.knn_graph(
            x[:, self._edge_definition._columns], self._edge_definition._nb_nearest_neighbours,
            batch, num_events=len(keep))
        return out
