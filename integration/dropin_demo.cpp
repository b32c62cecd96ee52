// Demo / acceptance driver for integration/engine_shims.cpp (built by `make -C oracle dropin`).
// usage: dropin_demo <directory> <mesh name>
// Runs the ENGINE'S OWN MeshLoader::Load (unmodified reference object code): it finds <dir>/<name>.obj, calls
// LoadOBJ -- which resolves to the CUDA-backed shim in this executable -- and then the engine's own
// WriteToE3DM writes <dir>/<name>.e3dm.  tests/ byte-compares that file with the pure reference's.
#include <cstdio>
#include <string>

#include "resources/geometry/geometry.hpp"
#include "systems/resource/loaders/mesh/mesh_loader.hpp"
#include "systems/resource/resources/mesh/mesh_resource.hpp"
#include "vkmesh_b200.h"

int main(int argc, char** argv) {
    if (argc < 3) {
        std::fprintf(stderr, "usage: %s <directory> <mesh name>\n", argv[0]);
        return 2;
    }
    if (vkm_device_count() == 0) {
        std::fprintf(stderr, "no CUDA device\n");
        return 3;
    }
    Engine::MeshLoader loader(0u, std::string(argv[1]));
    Engine::Resource* res = loader.Load(argv[2]);
    if (!res) {
        std::fprintf(stderr, "Load failed: %s\n", vkm_last_error());
        return 1;
    }
    auto* mesh = static_cast<Engine::MeshResource*>(res);
    for (auto& c : mesh->GetConfigs())
        std::printf("config '%s' material '%s': %u vertices, %u indices\n", c.name.c_str(), c.material_name.c_str(),
                    c.vertex_count, c.index_count);
    delete mesh;
    return 0;
}
