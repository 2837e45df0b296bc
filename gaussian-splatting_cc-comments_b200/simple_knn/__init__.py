"""Drop-in for the reference's `simple_knn` extension package (submodules/simple-knn): `from simple_knn._C import distCUDA2`
(scene/gaussian_model.py:20) resolves to this library's sm_100a kernels (csrc/gsr_knn.cu) through the C ABI."""
