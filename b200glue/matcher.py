"""`model.matcher.name: b200glue.matcher` -> resolved through __main_model__ like the
reference's own lightglue.py:666."""
import importlib

_m = importlib.import_module("cs566-project-lightglue_b200.matcher")
LightGlue = _m.LightGlue
__main_model__ = LightGlue
