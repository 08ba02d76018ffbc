"""Drop-in replacement of the reference's `simple_knn` package (KNN = submodules/simple-knn of
KTH-RPL/cloth-splatting): `from simple_knn._C import distCUDA2` (KNN/ext.cpp:15-17)."""
