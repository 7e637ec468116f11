"""Importable alias of the product package (whose directory name is not a Python identifier).
Lets the reference's registry select the matcher with `model.matcher.name: b200glue.matcher`
(gluefactory/models/__init__.py:7-30)."""
import importlib

_impl = importlib.import_module("cs566-project-lightglue_b200")
