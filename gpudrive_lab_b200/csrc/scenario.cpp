// See scenario.h. Compile with -ffp-contract=off: the running means and the polyline-area test
// are float expressions whose rounding decides which points survive.
#include "scenario.h"

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <thread>
#include <unordered_map>
#include <unordered_set>

#include "json_min.h"

namespace gd {

// ---------------------------------------------------------------------------------------------
// JSON -> ParsedMap
// ---------------------------------------------------------------------------------------------

static Vec2 readVec2(const JsonValue &j) {   // json_serialization.hpp:12-16
    return Vec2{j.at("x").asFloat(), j.at("y").asFloat()};
}

// from_json(MapObject), json_serialization.hpp:18-109
static void readObject(const JsonValue &j, ParsedObject &obj) {
    uint32_t i = 0;
    for (const auto &pos : j.at("position").arr) {
        if (i >= (uint32_t)MAX_POSITIONS) break;
        obj.position[i++] = readVec2(pos);
    }
    obj.numPositions = i;
    obj.width = j.at("width").asFloat();
    obj.length = j.at("length").asFloat();
    obj.height = j.at("height").asFloat();
    obj.id = (uint32_t)j.at("id").asInt();

    i = 0;
    for (const auto &h : j.at("heading").arr) {
        if (i >= (uint32_t)MAX_POSITIONS) break;
        obj.heading[i++] = h.asFloat();
    }
    obj.numHeadings = i;

    i = 0;
    for (const auto &v : j.at("velocity").arr) {
        if (i >= (uint32_t)MAX_POSITIONS) break;
        obj.velocity[i++] = readVec2(v);
    }
    obj.numVelocities = i;

    i = 0;
    for (const auto &v : j.at("valid").arr) {
        if (i >= (uint32_t)MAX_POSITIONS) break;
        obj.valid[i++] = v.asBool();
    }
    obj.numValid = i;

    obj.goal = readVec2(j.at("goalPosition"));
    const std::string &type = j.at("type").asString();
    if (type == "vehicle") obj.type = ET_Vehicle;
    else if (type == "pedestrian") obj.type = ET_Pedestrian;
    else if (type == "cyclist") obj.type = ET_Cyclist;
    else obj.type = ET_None;

    if (const JsonValue *m = j.find("mark_as_expert")) obj.markAsExpert = m->asBool();

    obj.meta[0] = obj.meta[1] = obj.meta[2] = obj.meta[3] = 0;
}

// from_json(MapRoad), json_serialization.hpp:111-244 (the per-road mean is never consumed; omitted)
static void readRoad(const JsonValue &j, ParsedRoad &road, float threshold) {
    const std::string &type = j.at("type").asString();
    if (type == "road_edge") road.type = ET_RoadEdge;
    else if (type == "road_line") road.type = ET_RoadLine;
    else if (type == "lane") road.type = ET_RoadLane;
    else if (type == "crosswalk") road.type = ET_CrossWalk;
    else if (type == "speed_bump") road.type = ET_SpeedBump;
    else if (type == "stop_sign") road.type = ET_StopSign;
    else road.type = ET_None;

    std::vector<Vec2> pts;
    const auto &geom = j.at("geometry").arr;
    pts.reserve(geom.size());
    for (const auto &p : geom) pts.push_back(readVec2(p));

    const int64_t num_segments = (int64_t)geom.size() - 1;
    const int64_t n = num_segments + 1;   // sample_every_n_ == 1
    road.geometry.clear();
    if (num_segments >= 10 &&
        (road.type == ET_RoadLane || road.type == ET_RoadEdge || road.type == ET_RoadLine)) {
        // Visvalingam-style reduction, json_serialization.hpp:144-203
        std::vector<char> skip((size_t)n, 0);
        bool changed = true;
        while (changed) {
            changed = false;
            int64_t k = 0;
            while (k < n - 1) {
                int64_t k1 = k + 1;
                while (k1 < n - 1 && skip[(size_t)k1]) k1++;
                if (k1 >= n - 1) break;
                int64_t k2 = k1 + 1;
                while (k2 < n && skip[(size_t)k2]) k2++;
                if (k2 >= n) break;
                const Vec2 p1 = pts[(size_t)k], p2 = pts[(size_t)k1], p3 = pts[(size_t)k2];
                // float expression, then 0.5 (double) * |.|, narrowed to float_t (:170)
                float cross = (p1.x - p3.x) * (p2.y - p1.y) - (p1.x - p2.x) * (p3.y - p1.y);
                float area = (float)(0.5 * (double)std::fabs(cross));
                if (area < threshold) {
                    skip[(size_t)k1] = 1;
                    k = k2;
                    changed = true;
                } else {
                    k = k1;
                }
            }
        }
        skip[0] = 0;
        skip[(size_t)n - 1] = 0;
        uint32_t kept = 0;
        for (int64_t k = 0; k < n; k++) {
            if (skip[(size_t)k]) continue;
            if (kept < (uint32_t)MAX_GEOMETRY) road.geometry.push_back(pts[(size_t)k]);
            kept++;
        }
        road.numPoints = kept;
    } else {
        for (int64_t k = 0; k < n && k < MAX_GEOMETRY; k++) road.geometry.push_back(pts[(size_t)k]);
        road.numPoints = (uint32_t)std::max<int64_t>(n, 0);
    }

    road.id = 0;
    if (const JsonValue *id = j.find("id")) road.id = (uint32_t)id->asInt();

    road.mapType = -1;   // MapType::UNKNOWN
    if (const JsonValue *me = j.find("map_element_id")) {
        int32_t v = (int32_t)me->asInt();
        road.mapType = (v == 4 || v >= 21 || v < -1) ? -1 : v;   // :224-231
    }
}

// calc_mean, json_serialization.hpp:246-279 (incremental float mean over valid object samples and
// every raw road point)
static Vec2 calcMean(const JsonValue &j) {
    float mx = 0, my = 0;
    int64_t cnt = 0;
    for (const auto &obj : j.at("objects").arr) {
        const auto &valid = obj.at("valid");
        size_t i = 0;
        for (const auto &pos : obj.at("position").arr) {
            const JsonValue &v = valid[i++];
            if (v.kind == JsonValue::Bool && v.b == false) continue;
            cnt++;
            float nx = pos.at("x").asFloat(), ny = pos.at("y").asFloat();
            mx += (nx - mx) / cnt;
            my += (ny - my) / cnt;
        }
    }
    for (const auto &road : j.at("roads").arr) {
        for (const auto &pt : road.at("geometry").arr) {
            cnt++;
            float nx = pt.at("x").asFloat(), ny = pt.at("y").asFloat();
            mx += (nx - mx) / cnt;
            my += (ny - my) / cnt;
        }
    }
    return Vec2{mx, my};
}

// from_json(Map), json_serialization.hpp:281-414
static ParsedMap readMap(const JsonValue &j, float threshold) {
    ParsedMap map;
    memset(map.mapName, 0, sizeof(map.mapName));
    memset(map.scenarioId, 0, sizeof(map.scenarioId));
    strncpy(map.mapName, j.at("name").asString().c_str(), sizeof(map.mapName));
    strncpy(map.scenarioId, j.at("scenario_id").asString().c_str(), sizeof(map.scenarioId));

    map.mean = calcMean(j);
    const auto &objs = j.at("objects").arr;
    const size_t numObjects = std::min(objs.size(), (size_t)MAX_OBJECTS);
    ParsedObject zero_obj;
    memset(&zero_obj, 0, sizeof(zero_obj));
    map.objects.assign(numObjects, zero_obj);

    const JsonValue &metadata = j.at("metadata");
    const int64_t sdc_index = metadata.at("sdc_track_index").asInt();

    std::unordered_map<int64_t, size_t> idToObjIdx;
    std::unordered_set<int64_t> ttp, ooi;
    size_t idx = 0;

    const auto &tracks = metadata.at("tracks_to_predict").arr;
    for (const auto &t : tracks) {
        int64_t ti = t.at("track_index").asInt();
        if (ti >= 0 && (size_t)ti < objs.size()) ttp.insert(ti);
        else fprintf(stderr, "Warning: Invalid track_index %lld in scene %s\n", (long long)ti,
                     j.at("name").asString().c_str());
    }
    for (const auto &o : metadata.at("objects_of_interest").arr) ooi.insert(o.asInt());

    auto difficultyOf = [&](int64_t track_index, int32_t *out) {
        for (const auto &t : tracks) {
            if (t.at("track_index").asInt() == track_index) {
                *out = (int32_t)t.at("difficulty").asInt();
                return;
            }
        }
    };

    // SDC first (:320-346)
    if (sdc_index >= 0 && (size_t)sdc_index < objs.size() && numObjects > 0) {
        readObject(objs[(size_t)sdc_index], map.objects[0]);
        map.objects[0].meta[0] = 1;
        int64_t sdc_id = (int32_t)map.objects[0].id;
        if (ttp.count(sdc_index)) {
            map.objects[0].meta[2] = 1;
            difficultyOf(sdc_index, &map.objects[0].meta[3]);
        }
        if (ooi.count(sdc_id)) map.objects[0].meta[1] = 1;
        idToObjIdx[sdc_id] = 0;
        idx = 1;
        ttp.erase(sdc_index);
        ooi.erase(sdc_id);
    }
    // tracks_to_predict (:349-373)
    for (size_t i = 0; i < objs.size() && idx < numObjects; i++) {
        if ((int64_t)i == sdc_index) continue;
        if (!ttp.count((int64_t)i)) continue;
        readObject(objs[i], map.objects[idx]);
        map.objects[idx].meta[2] = 1;
        difficultyOf((int64_t)i, &map.objects[idx].meta[3]);
        int64_t oid = (int32_t)map.objects[idx].id;
        if (ooi.count(oid)) {
            map.objects[idx].meta[1] = 1;
            ooi.erase(oid);
        }
        idToObjIdx[oid] = idx;
        idx++;
    }
    // objects_of_interest (:376-387)
    for (size_t i = 0; i < objs.size() && idx < numObjects; i++) {
        if ((int64_t)i == sdc_index) continue;
        int64_t oid = (int32_t)objs[i].at("id").asInt();
        if (!ooi.count(oid)) continue;
        readObject(objs[i], map.objects[idx]);
        map.objects[idx].meta[1] = 1;
        idToObjIdx[(int32_t)map.objects[idx].id] = idx;
        idx++;
    }
    // the rest (:390-399)
    for (size_t i = 0; i < objs.size() && idx < numObjects; i++) {
        if ((int64_t)i == sdc_index) continue;
        int64_t oid = (int32_t)objs[i].at("id").asInt();
        if (idToObjIdx.count(oid)) continue;
        readObject(objs[i], map.objects[idx]);
        idToObjIdx[(int32_t)map.objects[idx].id] = idx;
        idx++;
    }

    // roads (:402-413)
    const auto &roads = j.at("roads").arr;
    const size_t numRoads = std::min(roads.size(), (size_t)MAX_ROADS);
    map.roads.resize(numRoads);
    size_t segs = 0;
    for (size_t r = 0; r < numRoads; r++) {
        readRoad(roads[r], map.roads[r], threshold);
        segs += (map.roads[r].type <= ET_RoadLane) ? (size_t)(map.roads[r].numPoints - 1) : 1;
    }
    map.numRoadSegments = (uint32_t)segs;
    return map;
}

ParsedMap parseMapString(const char *begin, const char *end, float threshold) {
    JsonParser p(begin, end);
    JsonValue root = p.parse();
    return readMap(root, threshold);
}

ParsedMap parseMapFile(const std::string &path, float threshold) {
    std::ifstream in(path, std::ios::binary);
    if (!in.is_open()) throw std::runtime_error("cannot open scenario file: " + path);
    std::stringstream ss;
    ss << in.rdbuf();
    std::string text = ss.str();
    try {
        return parseMapString(text.data(), text.data() + text.size(), threshold);
    } catch (const std::exception &e) {
        throw std::runtime_error(path + ": " + e.what());
    }
}

// ---------------------------------------------------------------------------------------------
// Binary scenario pool
//   header : "GDPOOL1\0" | float threshold | uint32 count | uint64 offset[count + 1]
//   entry  : mapName[32] scenarioId[32] mean(2 f32) numRoadSegments(u32)
//            numObjects(u32) objects... numRoads(u32) roads...   (field by field, little endian, no padding)
// ---------------------------------------------------------------------------------------------
namespace {

const char kPoolMagic[8] = {'G', 'D', 'P', 'O', 'O', 'L', '1', '\0'};

struct BlobWriter {
    std::string out;
    void raw(const void *p, size_t n) { out.append(static_cast<const char *>(p), n); }
    template <typename T> void put(const T &v) { raw(&v, sizeof(T)); }
};

struct BlobReader {
    const char *p, *end;
    void raw(void *dst, size_t n) {
        if ((size_t)(end - p) < n) throw std::runtime_error("scenario pool entry is truncated");
        memcpy(dst, p, n); p += n;
    }
    template <typename T> T get() { T v; raw(&v, sizeof(T)); return v; }
};

std::string encodeMap(const ParsedMap &m) {
    BlobWriter w;
    w.raw(m.mapName, 32); w.raw(m.scenarioId, 32);
    w.put(m.mean.x); w.put(m.mean.y); w.put(m.numRoadSegments);
    w.put((uint32_t)m.objects.size());
    for (const ParsedObject &o : m.objects) {
        w.raw(o.position, sizeof(o.position)); w.put(o.length); w.put(o.width); w.put(o.height);
        w.raw(o.heading, sizeof(o.heading)); w.raw(o.velocity, sizeof(o.velocity));
        for (int i = 0; i < MAX_POSITIONS; i++) w.put((uint8_t)(o.valid[i] ? 1 : 0));
        w.put(o.goal.x); w.put(o.goal.y); w.put(o.type); w.raw(o.meta, sizeof(o.meta));
        w.put(o.numPositions); w.put(o.numHeadings); w.put(o.numVelocities); w.put(o.numValid);
        w.put(o.id); w.put((uint8_t)(o.markAsExpert ? 1 : 0));
    }
    w.put((uint32_t)m.roads.size());
    for (const ParsedRoad &r : m.roads) {
        w.put(r.id); w.put(r.mapType); w.put(r.type); w.put(r.numPoints);
        w.put((uint32_t)r.geometry.size());
        if (!r.geometry.empty()) w.raw(r.geometry.data(), sizeof(Vec2) * r.geometry.size());
    }
    return std::move(w.out);
}

ParsedMap decodeMap(const char *begin, const char *end) {
    BlobReader r{begin, end};
    ParsedMap m;
    r.raw(m.mapName, 32); r.raw(m.scenarioId, 32);
    m.mean.x = r.get<float>(); m.mean.y = r.get<float>(); m.numRoadSegments = r.get<uint32_t>();
    const uint32_t nobj = r.get<uint32_t>();
    if (nobj > (uint32_t)MAX_OBJECTS) throw std::runtime_error("scenario pool entry: object count out of range");
    m.objects.resize(nobj);
    for (ParsedObject &o : m.objects) {
        r.raw(o.position, sizeof(o.position)); o.length = r.get<float>(); o.width = r.get<float>(); o.height = r.get<float>();
        r.raw(o.heading, sizeof(o.heading)); r.raw(o.velocity, sizeof(o.velocity));
        for (int i = 0; i < MAX_POSITIONS; i++) o.valid[i] = r.get<uint8_t>() != 0;
        o.goal.x = r.get<float>(); o.goal.y = r.get<float>(); o.type = r.get<int32_t>(); r.raw(o.meta, sizeof(o.meta));
        o.numPositions = r.get<uint32_t>(); o.numHeadings = r.get<uint32_t>(); o.numVelocities = r.get<uint32_t>();
        o.numValid = r.get<uint32_t>(); o.id = r.get<uint32_t>(); o.markAsExpert = r.get<uint8_t>() != 0;
    }
    const uint32_t nroad = r.get<uint32_t>();
    if (nroad > (uint32_t)MAX_ROADS) throw std::runtime_error("scenario pool entry: road count out of range");
    m.roads.resize(nroad);
    for (ParsedRoad &rd : m.roads) {
        rd.id = r.get<uint32_t>(); rd.mapType = r.get<int32_t>(); rd.type = r.get<int32_t>(); rd.numPoints = r.get<uint32_t>();
        const uint32_t ng = r.get<uint32_t>();
        if (ng > (uint32_t)MAX_GEOMETRY) throw std::runtime_error("scenario pool entry: geometry count out of range");
        rd.geometry.resize(ng);
        if (ng) r.raw(rd.geometry.data(), sizeof(Vec2) * ng);
    }
    return m;
}

struct PoolHeader { float threshold; uint32_t count; std::vector<uint64_t> offsets; };

PoolHeader readPoolHeader(std::ifstream &in, const std::string &path) {
    char magic[8];
    PoolHeader h;
    in.read(magic, 8);
    if (!in || memcmp(magic, kPoolMagic, 8) != 0) throw std::runtime_error(path + ": not a scenario pool file");
    in.read(reinterpret_cast<char *>(&h.threshold), 4);
    in.read(reinterpret_cast<char *>(&h.count), 4);
    if (!in || h.count > (1u << 24)) throw std::runtime_error(path + ": bad scenario pool header");
    h.offsets.resize((size_t)h.count + 1);
    in.read(reinterpret_cast<char *>(h.offsets.data()), sizeof(uint64_t) * h.offsets.size());
    if (!in) throw std::runtime_error(path + ": truncated scenario pool header");
    return h;
}

// "<file>#<index>" with a non-empty all-digit index
bool splitPoolRef(const std::string &path, std::string *file, int *index) {
    const size_t k = path.rfind('#');
    if (k == std::string::npos || k + 1 >= path.size() || path.size() - k > 9) return false;
    int v = 0;
    for (size_t i = k + 1; i < path.size(); i++) {
        if (path[i] < '0' || path[i] > '9') return false;
        v = v * 10 + (path[i] - '0');
    }
    *file = path.substr(0, k); *index = v;
    return true;
}

} // namespace

void writeScenarioPool(const std::string &out_path, const std::vector<std::string> &json_paths, float threshold,
                       int threads) {
    const size_t n = json_paths.size();
    std::vector<std::string> blobs(n), errors(n);
    std::atomic<size_t> next{0};
    auto work = [&]() {
        for (size_t i = next.fetch_add(1); i < n; i = next.fetch_add(1)) {
            try { blobs[i] = encodeMap(loadParsedMap(json_paths[i], threshold)); }
            catch (const std::exception &e) { errors[i] = e.what(); }
        }
    };
    const int T = std::max(1, std::min<int>(threads > 0 ? threads : (int)std::thread::hardware_concurrency(), (int)std::max<size_t>(n, 1)));
    std::vector<std::thread> pool;
    for (int t = 1; t < T; t++) pool.emplace_back(work);
    work();
    for (auto &t : pool) t.join();
    for (size_t i = 0; i < n; i++) if (!errors[i].empty()) throw std::runtime_error(errors[i]);
    std::ofstream out(out_path, std::ios::binary | std::ios::trunc);
    if (!out.is_open()) throw std::runtime_error("cannot write scenario pool: " + out_path);
    const uint32_t count = (uint32_t)n;
    std::vector<uint64_t> offsets(n + 1);
    uint64_t off = 8 + 4 + 4 + sizeof(uint64_t) * (n + 1);
    for (size_t i = 0; i < n; i++) { offsets[i] = off; off += blobs[i].size(); }
    offsets[n] = off;
    out.write(kPoolMagic, 8);
    out.write(reinterpret_cast<const char *>(&threshold), 4);
    out.write(reinterpret_cast<const char *>(&count), 4);
    out.write(reinterpret_cast<const char *>(offsets.data()), sizeof(uint64_t) * offsets.size());
    for (const std::string &b : blobs) out.write(b.data(), (std::streamsize)b.size());
    if (!out) throw std::runtime_error("short write on scenario pool: " + out_path);
}

int scenarioPoolCount(const std::string &pool_path, float *threshold_out) {
    std::ifstream in(pool_path, std::ios::binary);
    if (!in.is_open()) throw std::runtime_error("cannot open scenario pool: " + pool_path);
    PoolHeader h = readPoolHeader(in, pool_path);
    if (threshold_out) *threshold_out = h.threshold;
    return (int)h.count;
}

ParsedMap readScenarioPoolEntry(const std::string &pool_path, int index, float threshold) {
    std::ifstream in(pool_path, std::ios::binary);
    if (!in.is_open()) throw std::runtime_error("cannot open scenario pool: " + pool_path);
    PoolHeader h = readPoolHeader(in, pool_path);
    if (index < 0 || (uint32_t)index >= h.count)
        throw std::runtime_error(pool_path + ": entry " + std::to_string(index) + " of " + std::to_string(h.count) + " does not exist");
    if (memcmp(&h.threshold, &threshold, 4) != 0)
        throw std::runtime_error(pool_path + ": built with polylineReductionThreshold " + std::to_string(h.threshold) +
                                 ", asked for " + std::to_string(threshold));
    const uint64_t a = h.offsets[index], b = h.offsets[index + 1];
    if (b < a || b - a > (1ull << 31)) throw std::runtime_error(pool_path + ": corrupt entry table");
    std::string blob((size_t)(b - a), '\0');
    in.seekg((std::streamoff)a);
    in.read(&blob[0], (std::streamsize)blob.size());
    if (!in) throw std::runtime_error(pool_path + ": truncated entry " + std::to_string(index));
    return decodeMap(blob.data(), blob.data() + blob.size());
}

ParsedMap loadParsedMap(const std::string &path, float threshold) {
    std::string file; int index = 0;
    if (splitPoolRef(path, &file, &index)) return readScenarioPoolEntry(file, index, threshold);
    return parseMapFile(path, threshold);
}

// ---------------------------------------------------------------------------------------------
// Reference math restated for the host (level_gen.cpp / dynamics.hpp / utils.hpp)
// ---------------------------------------------------------------------------------------------

// utils::quatToYaw, src/utils.hpp:20-25 with q.x = q.y = 0; `atan2` there binds to the double overload
float hostQuatToYaw(float qw, float qz) {
    float a = 2.0f * (qw * qz + 0.f * 0.f);
    float b = 1.0f - 2.0f * (0.f * 0.f + qz * qz);
    return (float)atan2((double)a, (double)b);
}

// Quat::angleAxis(angle, up), external/madrona/include/madrona/math.inl:680-691
void hostAngleAxisZ(float angle, float *qw, float *qz) {
    *qw = cosf(angle / 2.f);
    *qz = 1.f * sinf(angle / 2.f);
}

// utils::NormalizeAngle<float>, src/utils.hpp:11-14 (`fmod` binds to the double overload)
static float normalizeAngle(float angle) {
    const float ret = (float)fmod((double)angle, (double)kPiM2);
    return ret > kPi ? ret - kPiM2 : (ret < -kPi ? ret + kPiM2 : ret);
}

// inverseBicycleModel, src/dynamics.hpp:117-149
static void inverseBicycle(float yaw_q_w, float yaw_q_z, float vx, float vy,
                           float tq_w, float tq_z, float tvx, float tvy, float *out3) {
    const float dt = 0.1f;
    float speed = sqrtf(vx * vx + vy * vy + 0.f * 0.f);
    float target_speed = sqrtf(tvx * tvx + tvy * tvy + 0.f * 0.f);
    float accel = (target_speed - speed) / dt;
    float yaw = normalizeAngle(hostQuatToYaw(yaw_q_w, yaw_q_z));
    float target_yaw = normalizeAngle(hostQuatToYaw(tq_w, tq_z));
    target_yaw = atan2f(tvy, tvx);   // consts::useEstimatedYaw == true
    float denominator = (float)((double)(speed * dt) + 0.5 * (double)accel * (double)dt * (double)dt);
    float steering = 0.f;
    if (denominator != 0) steering = (target_yaw - yaw) / denominator;
    out3[0] = accel;
    out3[1] = steering;
    out3[2] = 0.f;
}

// inverseDeltaModel, src/dynamics.hpp:151-184
static void inverseDelta(float qw, float qz, float px, float py, float tqw, float tqz,
                         float tpx, float tpy, float *out3) {
    float yaw = hostQuatToYaw(qw, qz);
    float target_yaw = hostQuatToYaw(tqw, tqz);
    float dx = tpx - px, dy = tpy - py, dyaw = target_yaw - yaw;
    dx = fmaxf(-6.0f, fminf(dx, 6.0f));
    dy = fmaxf(-6.0f, fminf(dy, 6.0f));
    float c = std::cos(-yaw), s = std::sin(-yaw);
    float ldx = dx * c - dy * s;
    float ldy = dx * s + dy * c;
    out3[0] = fmaxf(-6.0f, fminf(ldx, 6.0f));
    out3[1] = fmaxf(-6.0f, fminf(ldy, 6.0f));
    out3[2] = normalizeAngle(dyaw);
}

// getZeroAction, src/level_gen.hpp:16-38
static void zeroAction(int model, float *a10) {
    for (int k = 0; k < kActionSize; k++) a10[k] = 0.f;
    if (model == DM_State) a10[2] = 1.f;   // position {0,0,1}
}

// ---------------------------------------------------------------------------------------------
// ParsedMap -> SceneData  (createPersistentEntities, src/level_gen.cpp:396-465)
// ---------------------------------------------------------------------------------------------

static void pushRoad(SceneData &s, float x, float y, float z, float angle, float d0, float d1,
                     float d2, int32_t type, uint32_t id, int32_t mapType) {
    float qw, qz;
    hostAngleAxisZ(angle, &qw, &qz);
    s.rx.push_back(x); s.ry.push_back(y); s.rz.push_back(z);
    s.rqw.push_back(qw); s.rqz.push_back(qz);
    s.rd0.push_back(d0); s.rd1.push_back(d1); s.rd2.push_back(d2);
    s.ryaw.push_back(hostQuatToYaw(qw, qz));       // setRoadEntitiesProps, level_gen.hpp:40-65
    s.rtype.push_back(type);
    s.rid.push_back((float)id);
    s.rmaptype.push_back(mapType);
    s.numRoads++;
}

SceneData buildScene(const ParsedMap &map, const Params &params, int max_agents,
                     const int32_t *deleted, int num_deleted) {
    SceneData s;
    for (int i = 0; i < 32; i++) {   // level_gen.cpp:400-408 (char -> char32_t)
        s.mapName[i] = (int32_t)(char32_t)map.mapName[i];
        s.scenarioId[i] = (int32_t)(char32_t)map.scenarioId[i];
    }
    s.mean[0] = map.mean.x; s.mean[1] = map.mean.y; s.mean[2] = 0.f;
    const float mx = map.mean.x, my = map.mean.y;
    const int model = params.dynamicsModel;

    auto isDeleted = [&](uint32_t id) {
        for (int i = 0; i < num_deleted; i++)
            if (deleted[i] == (int32_t)id) return true;
        return false;
    };

    // ---- agents (level_gen.cpp:421-434, createAgent :131-164) ----
    for (size_t oc = 0; oc < map.objects.size() && s.numAgents < max_agents; oc++) {
        const ParsedObject &o = map.objects[oc];
        // shouldAgentBeCreated, :353-394
        if (params.readFromTracksToPredict) {
            if (isDeleted(o.id)) continue;
        } else {
            if (params.ignoreNonVehicles && (o.type == ET_Pedestrian || o.type == ET_Cyclist)) continue;
            if (params.initOnlyValidAgentsAtFirstStep && !o.valid[0]) continue;
            if (isDeleted(o.id)) continue;
        }

        const size_t a = (size_t)s.numAgents;
        s.traj.resize((a + 1) * kTrajSize, 0.f);
        float *tr = &s.traj[a * kTrajSize];
        // Trajectory component starts zeroed (fresh ECS row), then populateExpertTrajectory :56-100
        for (int i = 0; i < kTrajLen; i++) zeroAction(DM_Classic, tr + kTrajInv + i * kActionSize);
        const int np = (int)o.numPositions;
        for (int i = 0; i < np; i++) {
            tr[kTrajPos + 2 * i] = o.position[i].x - mx;
            tr[kTrajPos + 2 * i + 1] = o.position[i].y - my;
            tr[kTrajVel + 2 * i] = o.velocity[i].x;
            tr[kTrajVel + 2 * i + 1] = o.velocity[i].y;
            tr[kTrajHead + i] = o.heading[i];
            tr[kTrajValid + i] = (float)o.valid[i];
            zeroAction(model, tr + kTrajInv + i * kActionSize);
        }
        if (model == DM_InvertibleBicycle || model == DM_DeltaLocal) {
            for (int i = np - 2; i >= 0; i--) {
                // (the zero action written for invalid spans at :72-75 is overwritten below, as in the reference)
                float qw, qz, tqw, tqz;
                hostAngleAxisZ(tr[kTrajHead + i], &qw, &qz);
                hostAngleAxisZ(tr[kTrajHead + i + 1], &tqw, &tqz);
                float *inv = tr + kTrajInv + i * kActionSize;
                if (model == DM_InvertibleBicycle) {
                    inverseBicycle(qw, qz, tr[kTrajVel + 2 * i], tr[kTrajVel + 2 * i + 1], tqw, tqz,
                                   tr[kTrajVel + 2 * (i + 1)], tr[kTrajVel + 2 * (i + 1) + 1], inv);
                } else {
                    inverseDelta(qw, qz, tr[kTrajPos + 2 * i], tr[kTrajPos + 2 * i + 1], tqw, tqz,
                                 tr[kTrajPos + 2 * (i + 1)], tr[kTrajPos + 2 * (i + 1) + 1], inv);
                }
            }
        }

        s.vsize.push_back(o.length); s.vsize.push_back(o.width); s.vsize.push_back(o.height);
        // Scale = {len/2, wid/2, 1} * vehicleLengthScale (:140-141)
        s.scale.push_back((o.length / 2) * kVehicleLengthScale);
        s.scale.push_back((o.width / 2) * kVehicleLengthScale);
        s.scale.push_back(1.f * kVehicleLengthScale);
        const float gx = o.goal.x - mx, gy = o.goal.y - my;
        s.goal.push_back(gx); s.goal.push_back(gy);
        s.type.push_back(o.type);
        s.id.push_back((int32_t)o.id);

        // isAgentStatic :102-113. MetaData is read before it is assigned (:106 vs :154); a fresh
        // component row is zero, so isTrackToPredict != -1 holds for every agent in that mode.
        bool is_static;
        if (params.readFromTracksToPredict) {
            is_static = false;
        } else {
            float ddx = gx - tr[kTrajPos + 0], ddy = gy - tr[kTrajPos + 1];
            bool close = sqrtf(ddx * ddx + ddy * ddy) < kStaticThreshold;
            is_static = !params.isStaticAgentControlled && close;
        }
        const int32_t resp = is_static ? RT_Static : RT_Dynamic;
        s.response.push_back(resp);

        // isAgentControllable :115-129
        bool ctrl;
        if (params.readFromTracksToPredict) {
            ctrl = (uint32_t)s.numControlled < params.maxNumControlledAgents;
        } else {
            ctrl = (uint32_t)s.numControlled < params.maxNumControlledAgents &&
                   tr[kTrajValid + 0] != 0.f && resp == RT_Dynamic && !o.markAsExpert;
        }
        s.controlled.push_back(ctrl ? 1 : 0);
        s.numControlled += ctrl ? 1 : 0;
        for (int k = 0; k < 4; k++) s.meta.push_back(o.meta[k]);
        s.numAgents++;
    }

    // ---- road entities (createRoadEntities :258-296) ----
    for (size_t rc = 0; rc < map.roads.size() && s.numRoads < kMaxRoadEntityCount; rc++) {
        const ParsedRoad &r = map.roads[rc];
        const size_t have = r.geometry.size();
        switch (r.type) {
        case ET_RoadEdge: case ET_RoadLine: case ET_RoadLane: {
            // makeRoadEdge :166-185
            const float z = 1 + (r.type == ET_RoadEdge ? kLidarRoadEdgeOffset : kLidarRoadLineOffset);
            const size_t npts = std::min((size_t)r.numPoints, have);   // reference reads past the array; clamp
            for (size_t j = 1; j + 1 <= npts && npts >= 1; j++) {
                const Vec2 p1 = r.geometry[j - 1], p2 = r.geometry[j];
                const float sx = p1.x - mx, sy = p1.y - my, ex = p2.x - mx, ey = p2.y - my;
                const float px = (sx + ex) / 2, py = (sy + ey) / 2;
                const float ang = (float)atan2((double)(ey - sy), (double)(ex - sx));
                const float ddx = sx - ex, ddy = sy - ey, ddz = z - z;
                const float len = sqrtf(ddx * ddx + ddy * ddy + ddz * ddz);
                pushRoad(s, px, py, z, ang, len / 2, (float)0.1, (float)0.1, r.type, r.id, r.mapType);
                if (s.numRoads >= kMaxRoadEntityCount) break;
            }
        } break;
        case ET_CrossWalk: case ET_SpeedBump: {
            // makeCube :191-241
            if (have < 4) break;
            Vec2 pts[4] = {r.geometry[0], r.geometry[1], r.geometry[2], r.geometry[3]};
            float lengths[4];
            for (int i = 0; i < 4; i++) {
                const Vec2 &a = pts[i], &b = pts[(i + 1) % 4];
                lengths[i] = (float)sqrt(pow((double)(b.x - a.x), 2) + pow((double)(b.y - a.y), 2));
            }
            int maxi = 0, mini = 0;
            for (int i = 1; i < 4; i++) {
                if (lengths[i] > lengths[maxi]) maxi = i;
                if (lengths[i] < lengths[mini]) mini = i;
            }
            const Vec2 &st = pts[maxi], &en = pts[(maxi + 1) % 4];
            const float ang = (float)atan2((double)(en.y - st.y), (double)(en.x - st.x));
            float sumx = 0.f, sumy = 0.f;
            for (const Vec2 &p : pts) { sumx += p.x; sumy += p.y; }
            pushRoad(s, sumx / 4 - mx, sumy / 4 - my, 1 + kLidarRoadLineOffset, ang,
                     lengths[maxi] / 2, lengths[mini] / 2, (float)0.1, r.type, r.id, r.mapType);
        } break;
        case ET_StopSign: {
            // makeStopSign :243-256
            if (have < 1) break;
            pushRoad(s, r.geometry[0].x - mx, r.geometry[0].y - my, 1.f, 0.f, (float)0.2, (float)0.2,
                     1.f, ET_StopSign, r.id, r.mapType);
        } break;
        default: break;
        }
    }
    return s;
}

} // namespace gd
