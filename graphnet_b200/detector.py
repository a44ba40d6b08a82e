"""Per-feature standardisation of pulse maps (host side, elementwise).

Mirrors the reference's `Detector._standardize` contract
(/root/reference/src/graphnet/models/detector/detector.py:71-90): every input
column is mapped through the function registered under its feature name; an
unknown feature name is an error (KeyError), as in the reference.

Feature maps follow models/detector/icecube.py:21-48 (IceCube86),
:126-170 (IceCubeUpgrade) and models/detector/prometheus.py:21-38
(ORCA150SuperDense / Prometheus).
"""

from __future__ import annotations

from typing import Callable, Dict, List, Optional

import torch
from torch import Tensor


class Detector:
    """Base class: column-wise standardisation by feature name."""

    xyz: List[str] = []

    def __init__(self, replace_with_identity: Optional[List[str]] = None) -> None:
        self._replace_with_identity = replace_with_identity

    def feature_map(self) -> Dict[str, Callable[[Tensor], Tensor]]:
        raise NotImplementedError

    def __call__(self, input_features: Tensor, input_feature_names: List[str]) -> Tensor:
        return self.forward(input_features, input_feature_names)

    def forward(self, input_features: Tensor, input_feature_names: List[str]) -> Tensor:
        """detector.py:36-40: standardise a [n, d] tensor in place and return it."""
        return self._standardize(input_features, input_feature_names)

    def _standardize(self, input_features: Tensor, input_feature_names: List[str]) -> Tensor:
        fmap = self.feature_map()
        if self._replace_with_identity is not None:
            for f in self._replace_with_identity:
                fmap[f] = self._identity
        for idx, feature in enumerate(input_feature_names):
            if feature not in fmap:
                raise KeyError(f"No Standardization function found for '{feature}'")
            input_features[:, idx] = fmap[feature](input_features[:, idx])
        return input_features

    @staticmethod
    def _identity(x: Tensor) -> Tensor:
        return x


class IceCube86(Detector):
    """icecube.py:21-48."""

    xyz = ["dom_x", "dom_y", "dom_z"]

    def feature_map(self) -> Dict[str, Callable[[Tensor], Tensor]]:
        return {
            "dom_x": lambda x: x / 500.0,
            "dom_y": lambda x: x / 500.0,
            "dom_z": lambda x: x / 500.0,
            "dom_time": lambda x: (x - 1.0e04) / 3.0e4,
            "charge": torch.log10,
            "rde": lambda x: (x - 1.25) / 0.25,
            "pmt_area": lambda x: x / 0.05,
            "hlc": self._identity,
        }


class IceCubeUpgrade(Detector):
    """icecube.py:126-170."""

    xyz = ["dom_x", "dom_y", "dom_z"]

    def feature_map(self) -> Dict[str, Callable[[Tensor], Tensor]]:
        return {
            "dom_x": lambda x: x / 500.0,
            "dom_y": lambda x: x / 500.0,
            "dom_z": lambda x: x / 500.0,
            "dom_time": lambda x: (x / 2e04) - 1.0,
            "charge": lambda x: torch.log10(x) / 2.0,
            "rde": self._identity,
            "pmt_area": lambda x: x / 0.05,
            "string": lambda x: (x - 50.0) / 50.0,
            "pmt_number": lambda x: x / 20.0,
            "dom_number": lambda x: (x - 60.0) / 60.0,
            "pmt_dir_x": self._identity,
            "pmt_dir_y": self._identity,
            "pmt_dir_z": self._identity,
            "dom_type": lambda x: x / 130.0,
            "hlc": self._identity,
        }


class Prometheus(Detector):
    """prometheus.py:21-38 (ORCA150SuperDense, aliased as Prometheus :365)."""

    xyz = ["sensor_pos_x", "sensor_pos_y", "sensor_pos_z"]

    def feature_map(self) -> Dict[str, Callable[[Tensor], Tensor]]:
        return {
            "sensor_pos_x": lambda x: x / 100,
            "sensor_pos_y": lambda x: x / 100,
            "sensor_pos_z": lambda x: (x + 350) / 100,
            "t": lambda x: x / 1.05e04,
        }


ORCA150SuperDense = Prometheus

# Feature name lists, data/constants.py:7-25
FEATURES_ICECUBE86 = ["dom_x", "dom_y", "dom_z", "dom_time", "charge", "rde", "pmt_area"]
FEATURES_UPGRADE = FEATURES_ICECUBE86 + [
    "string", "pmt_number", "dom_number", "pmt_dir_x", "pmt_dir_y", "pmt_dir_z", "dom_type",
]
FEATURES_PROMETHEUS = ["sensor_pos_x", "sensor_pos_y", "sensor_pos_z", "t"]
