// Host-side scenario loader: WOMD-shaped JSON -> ParsedMap -> SceneData (flat SoA arrays that
// the engine uploads into the device scenario pool). Replaces MapReader + from_json(Map)
// (src/MapReader.cpp:48-62, src/json_serialization.hpp:12-414) and the entity construction of
// createPersistentEntities (src/level_gen.cpp:56-465). Runs once per scenario, never per step.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "gd_types.h"

namespace gd {

struct Vec2 { float x, y; };

// == MapObject, src/init.hpp:22-43
struct ParsedObject {
    Vec2 position[MAX_POSITIONS];
    float length, width, height;
    float heading[MAX_POSITIONS];
    Vec2 velocity[MAX_POSITIONS];
    bool valid[MAX_POSITIONS];
    Vec2 goal;
    int32_t type;
    int32_t meta[4];      // isSdc, isObjectOfInterest, isTrackToPredict, difficulty
    uint32_t numPositions, numHeadings, numVelocities, numValid;
    uint32_t id;
    bool markAsExpert;
};

// == MapRoad, src/init.hpp:45-54
struct ParsedRoad {
    std::vector<Vec2> geometry;   // at most MAX_GEOMETRY points kept
    uint32_t id;
    int32_t mapType;
    int32_t type;
    uint32_t numPoints;           // may exceed geometry.size() like the reference (json_serialization.hpp:203)
};

// == Map, src/init.hpp:56-73
struct ParsedMap {
    std::vector<ParsedObject> objects;   // numObjects entries, reference ordering
    std::vector<ParsedRoad> roads;
    uint32_t numRoadSegments;
    Vec2 mean;
    char mapName[32];
    char scenarioId[32];
};

// One scenario after entity construction for a given Params / max_agents / deleted-agent list.
struct SceneData {
    int32_t numAgents = 0;
    int32_t numRoads = 0;
    int32_t numControlled = 0;
    float mean[3] = {0, 0, 0};
    int32_t mapName[32];
    int32_t scenarioId[32];

    // agents, creation order (== export slot order)
    std::vector<float> traj;        // [numAgents][1456]  Trajectory rows
    std::vector<float> vsize;       // [numAgents][3]     VehicleSize length,width,height
    std::vector<float> scale;       // [numAgents][3]     Scale d0,d1,d2
    std::vector<float> goal;        // [numAgents][2]
    std::vector<int32_t> type;      // EntityType
    std::vector<int32_t> id;        // AgentID
    std::vector<int32_t> response;  // ResponseType
    std::vector<int32_t> controlled;
    std::vector<int32_t> meta;      // [numAgents][4]

    // road entities, creation order
    std::vector<float> rx, ry, rz;      // Position
    std::vector<float> rqw, rqz;        // Rotation (angleAxis about +z: x = y = 0)
    std::vector<float> rd0, rd1, rd2;   // Scale
    std::vector<float> ryaw;            // quatToYaw(rot) = MapObservation.heading
    std::vector<int32_t> rtype;         // EntityType
    std::vector<float> rid;             // float(RoadMapId)
    std::vector<int32_t> rmaptype;      // MapType
};

// Throws std::runtime_error on malformed input.
ParsedMap parseMapFile(const std::string &path, float polylineReductionThreshold);
ParsedMap parseMapString(const char *begin, const char *end, float polylineReductionThreshold);

// Binary scenario pool: the ParsedMaps of many scenarios in one file, so that a pool is
// parsed from JSON once (in parallel) and afterwards read back at memory speed. The file depends on the
// polyline reduction threshold only (it is applied while parsing, MapReader / json_serialization.hpp:160-240);
// every other parameter still acts in buildScene. A scene path of the form "<file>#<index>" names entry
// <index> of pool <file> wherever a JSON path is accepted.
void writeScenarioPool(const std::string &out_path, const std::vector<std::string> &json_paths, float threshold,
                       int threads);
int scenarioPoolCount(const std::string &pool_path, float *threshold_out);
ParsedMap readScenarioPoolEntry(const std::string &pool_path, int index, float threshold);
// JSON file or pool reference
ParsedMap loadParsedMap(const std::string &path, float threshold);

// deleted: agent ids excluded from creation (DeletedAgents singleton), may be null.
SceneData buildScene(const ParsedMap &map, const Params &params, int max_agents,
                     const int32_t *deleted, int num_deleted);

// Reference helpers restated for host use (bit-for-bit the same glibc calls as the g++ build
// of the reference makes: double atan2 / fmod, float sinf / cosf).
float hostQuatToYaw(float qw, float qz);
void hostAngleAxisZ(float angle, float *qw, float *qz);

} // namespace gd
