"""B200-native EnergyEquivGNN hot path (drop-in for gnn/model.py of igrega348/energy-equiv-lattice-gnn)."""
from .data import Batch, LatticeData, collate, synthetic_batch, synthetic_lattice  # noqa: F401
from .model import EnergyEquivGNN  # noqa: F401

__all__ = ["EnergyEquivGNN", "Batch", "LatticeData", "collate", "synthetic_batch", "synthetic_lattice"]
