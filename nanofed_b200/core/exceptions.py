"""Error hierarchy shared by every layer.

Parity: reference ``nanofed/core/exceptions.py:1-16`` (same three names, same
inheritance), so ``except NanoFedError`` written against the reference keeps
working.
"""


class NanoFedError(Exception):
    """Root of all framework errors."""


class AggregationError(NanoFedError):
    """Client updates could not be combined into a global model."""


class ModelManagerError(NanoFedError):
    """A checkpoint could not be written, found or restored."""
