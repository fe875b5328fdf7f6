"""Layer D: flat layouts, arenas, symmetric peer memory, rank bootstrap."""

from .flat import FlatEntry, FlatLayout

__all__ = ["FlatEntry", "FlatLayout"]
