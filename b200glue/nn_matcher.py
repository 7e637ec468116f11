"""`model.matcher.name: b200glue.nn_matcher` -> the B200 NearestNeighborMatcher (resolved through __main_model__,
gluefactory/models/__init__.py:20-25)."""
import importlib

_m = importlib.import_module("cs566-project-lightglue_b200.nn_matcher")
NearestNeighborMatcher = _m.NearestNeighborMatcher
__main_model__ = NearestNeighborMatcher
