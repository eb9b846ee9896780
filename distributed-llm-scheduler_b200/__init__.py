# Package directory for `dls_b200` (see ../dls_b200/__init__.py, which puts this directory on its __path__).
