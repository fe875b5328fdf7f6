"""Layer D: flat layouts, arenas, symmetric peer memory, rank bootstrap."""

from .arena import Arena
from .flat import FlatEntry, FlatLayout

__all__ = ["Arena", "FlatEntry", "FlatLayout"]
