# Supplied by baseline/install_ref.py: the reference tree has no asgd/__init__.py.
