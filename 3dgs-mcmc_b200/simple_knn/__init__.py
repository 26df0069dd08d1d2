"""Drop-in for the reference's ``simple_knn`` extension package (submodules/simple-knn): ``simple_knn._C.distCUDA2``,
backed by lib3dgs_mcmc_b200.so (csrc/knn.cu). With ``3dgs-mcmc_b200/`` on ``sys.path``,
``from simple_knn._C import distCUDA2`` (scene/gaussian_model_static.py:24) resolves here."""
