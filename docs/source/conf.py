"""Sphinx configuration (``make docs``).  The narrative pages are the Markdown files under
``docs/`` and the repository root; this tree only adds the API reference and the index."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))

project = "nanofed_b200"
author = "nanofed_b200 developers"
extensions = ["sphinx.ext.autodoc", "sphinx.ext.napoleon", "sphinx.ext.viewcode"]
autodoc_mock_imports = ["torch", "aiohttp", "pydantic", "cryptography", "numpy"]
autodoc_member_order = "bysource"
templates_path = []
exclude_patterns = []
html_theme = "alabaster"
