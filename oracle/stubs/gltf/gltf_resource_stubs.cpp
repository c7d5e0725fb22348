// TEST INFRASTRUCTURE — no-op bodies for the one GPU resource class the reference's glTF loader constructs through
// the reference's own header (include/resource_mesh.h).  No graphics API is touched.
#include "resource_mesh.h"

namespace engine {
Mesh::Mesh(GFXDevice* gfx, const std::vector<Vertex>& vertices, const std::vector<uint32_t>& indices)
    : m_gfx(gfx), m_vb(nullptr), m_ib(nullptr), m_count(static_cast<uint32_t>(indices.size()))
{
    (void)vertices;
}
Mesh::~Mesh() {}
}  // namespace engine
