"""Name shim: `import simple_knn` -> splatloc_b200.simple_knn."""
