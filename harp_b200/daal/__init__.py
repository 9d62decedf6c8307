"""Harp-DAAL equivalents: the ``harp-daal-interface`` support layer (``comm``, ``datasource``,
``initialize``) and the ``ml/daal`` algorithms (``linalg``, ``classify``, ``models``)."""
