from splatloc_b200._C import distCUDA2  # noqa: F401
