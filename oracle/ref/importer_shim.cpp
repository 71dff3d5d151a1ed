// Stand-in for external/madrona/src/importer/importer.cpp (ImportedAssets::importFromDisk).
// The stock file does not compile with g++ (importer.cpp:33, brace-init through an
// explicit constructor) and drags in the glTF/USD loaders + simdjson. The reference
// only ever imports two .obj collision hulls (src/mgr.cpp:269-290), so this glue
// dispatches straight to the reference's own OBJLoader (importer/obj.cpp, compiled
// unmodified). No arithmetic lives here. Test infrastructure (oracle/_ref) only.
#include <madrona/importer.hpp>
#include <madrona/dyn_array.hpp>
#include <string_view>
#include "importer/obj.hpp"

namespace madrona::imp {

Optional<ImportedAssets> ImportedAssets::importFromDisk(
    Span<const char * const> paths, Span<char> err_buf, bool)
{
    ImportedAssets imported {
        .geoData = GeometryData {
            .positionArrays = DynArray<DynArray<math::Vector3>>(0),
            .normalArrays = DynArray<DynArray<math::Vector3>>(0),
            .tangentAndSignArrays = DynArray<DynArray<math::Vector4>>(0),
            .uvArrays = DynArray<DynArray<math::Vector2>>(0),
            .indexArrays = DynArray<DynArray<uint32_t>>(0),
            .faceCountArrays = DynArray<DynArray<uint32_t>>(0),
            .meshArrays = DynArray<DynArray<SourceMesh>>(0),
        },
        .objects = DynArray<SourceObject>(0),
        .materials = DynArray<SourceMaterial>(0),
        .instances = DynArray<SourceInstance>(0),
    };

    OBJLoader obj_loader(err_buf);
    for (const char *path : paths) {
        std::string_view v(path);
        auto dot = v.rfind('.');
        if (dot == v.npos || v.substr(dot + 1) != "obj") {
            snprintf(err_buf.data(), err_buf.size(), "gd_ref: only .obj supported (%s)", path);
            return Optional<ImportedAssets>::none();
        }
        if (!obj_loader.load(path, imported)) {
            return Optional<ImportedAssets>::none();
        }
    }
    return imported;
}

}
