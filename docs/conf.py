# Sphinx configuration (MyST markdown + autodoc).
import os
import sys

sys.path.insert(0, os.path.abspath(".."))
project = "Megatron-LLM for B200"
author = "the authors"
extensions = ["sphinx.ext.autodoc", "sphinx.ext.autosummary", "sphinx.ext.napoleon", "sphinx.ext.viewcode",
              "myst_parser"]
templates_path = ["_templates"]
autosummary_generate = True        # pages from _templates/autosummary/{module,class,base}.rst
source_suffix = {".rst": "restructuredtext", ".md": "markdown"}
autodoc_mock_imports = ["torch", "flash_attn", "sentencepiece", "transformers", "wandb", "faiss"]
html_theme = "alabaster"
exclude_patterns = ["_build"]
