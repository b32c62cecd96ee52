// TEST INFRASTRUCTURE ONLY (oracle).  systems/geometry/geometry_system.cpp is compiled into oracle/_ref for its two
// closed-form generators (GeneratePlainConfig / GenerateCubeConfig, geometry_system.cpp:182-439).  Its other
// member functions reference the material system and the renderer front end, which are out of scope and not
// built; these definitions only satisfy the linker and are never reached by the oracle.
#include <cstdlib>

#include "renderer/renderer.hpp"
#include "systems/material/material_system.hpp"

namespace Engine {

MaterialSystem* MaterialSystem::instance = nullptr;
Material* MaterialSystem::AcquireMaterial(std::string, b8) { std::abort(); }
RendererFrontend* RendererFrontend::GetInstance() { std::abort(); }
Geometry* RendererFrontend::CreateGeometry(GeometryCreateInfo&) { std::abort(); }

}  // namespace Engine
