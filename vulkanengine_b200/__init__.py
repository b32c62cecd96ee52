"""B200-native mesh-import geometry path (expand -> dedup -> AABB -> tangents/normals).

The product is the C-ABI shared library built from ``vulkanengine_b200/csrc`` (declared in
``include/vkmesh_b200.h``); this Python package only holds the build recipe, ctypes bindings
for tests / bench, and the synthetic mesh generators.
"""
