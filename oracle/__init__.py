"""Independent NumPy float64 restatements of the reference's hot loops, written from the Java sources and sharing no code
with harp_b200: the correctness oracles of the CPU tests (``tests/test_oracles.py``, ``tests/test_collectives_mp.py``).

``kmeans``  Lloyd iteration (CenCalcTask / CenMergeTask)          ``mf``     SGD pass (SGDMPTask) and CCD++ (CCDMPTask)
``lda``     collapsed Gibbs pass + likelihood (LDAMPTask, CalcLikelihoodTask)
``stats``   covariance and normal-equation partials (COVDaalCollectiveMapper, LinRegDaalCollectiveMapper)
"""
