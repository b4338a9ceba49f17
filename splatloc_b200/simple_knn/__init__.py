"""`simple_knn`-compatible front end: `from splatloc_b200.simple_knn._C import distCUDA2`."""
