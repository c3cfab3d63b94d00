"""shim: pytorch3d (see ../README.md)"""
