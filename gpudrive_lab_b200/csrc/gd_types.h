// Shared POD types and constants of the B200 step engine (host + device).
// Values and layouts follow the reference: src/consts.hpp, src/types.hpp, src/init.hpp.
#pragma once
#include <cstdint>

namespace gd {

// src/consts.hpp:11-13 (kMaxAgentCount is a ctor constant here: GdConfig.max_agents)
constexpr int kMaxRoadEntityCount = 10000;
constexpr int kMaxAgentMapObs = 200;        // K
constexpr int kEpisodeLen = 91;             // src/consts.hpp:34
constexpr int kTrajLen = 91;                // src/consts.hpp:57
constexpr int kNumLidarSamples = 50;        // src/consts.hpp:37
constexpr float kStaticThreshold = 0.2f;    // src/consts.hpp:17
constexpr float kVehicleLengthScale = 0.7f; // src/consts.hpp:25
constexpr float kLidarCarOffset = 0.5f;     // src/consts.hpp:42-46
constexpr float kLidarRoadEdgeOffset = 0.1f;
constexpr float kLidarRoadLineOffset = -0.1f;
constexpr float kLidarDistance = 200.f;
constexpr float kPi = 3.14159265358979323846f;  // madrona::math::pi (float)
constexpr float kPiM2 = kPi * 2.f;              // madrona::math::pi_m2
constexpr float kLidarAngle = kPi / 3;
constexpr float kPadX = -11000.f;           // consts::kPaddingPosition, src/consts.hpp:61-62
constexpr float kPadY = -11000.f;
constexpr float kPadZ = 3.402823466e+38f;   // FLT_MAX
// physics expansion (external/madrona/src/physics/physics.cpp:108-111, deltaT = 0.04)
constexpr float kLeafVelExpansion = 2.f * 0.04f;
constexpr float kLeafAccelExpansion = 100.f * 0.04f * 0.04f; // 9.8*... see physics.cpp; evaluated there as 0.16

// JSON loader caps, src/init.hpp:9-12
constexpr int MAX_OBJECTS = 515;
constexpr int MAX_ROADS = 956;
constexpr int MAX_POSITIONS = 91;
constexpr int MAX_GEOMETRY = 1746;

// src/types.hpp:24-38
enum EntityType : int32_t {
    ET_None = 0, ET_RoadEdge, ET_RoadLine, ET_RoadLane, ET_CrossWalk, ET_SpeedBump,
    ET_StopSign, ET_Vehicle, ET_Pedestrian, ET_Cyclist, ET_Padding, ET_NumTypes
};
// madrona::phys::ResponseType (external/madrona/include/madrona/physics.hpp)
enum ResponseType : int32_t { RT_Dynamic = 0, RT_Kinematic = 1, RT_Static = 2 };

// src/init.hpp:76-109
enum RewardType : int32_t { RW_DistanceBased = 0, RW_OnGoalAchieved = 1, RW_Dense = 2 };
enum CollisionBehaviour : int32_t { CB_AgentStop = 0, CB_AgentRemoved = 1, CB_Ignore = 2 };
enum DynamicsModel : int32_t { DM_Classic = 0, DM_InvertibleBicycle = 1, DM_DeltaLocal = 2, DM_State = 3 };
enum RoadObsAlgo : int32_t { RO_KNearest = 0, RO_AllEntities = 1 };

// Export row sizes (src/types.hpp *ExportSize)
constexpr int kActionSize = 10;
constexpr int kInfoSize = 5;
constexpr int kSelfObsSize = 8;
constexpr int kMapObsSize = 9;
constexpr int kPartnerObsSize = 9;
constexpr int kLidarRowSize = 4;
constexpr int kAbsObsSize = 14;
constexpr int kTrajSize = 2 * 2 * kTrajLen + 2 * kTrajLen + kActionSize * kTrajLen; // 1456
constexpr int kMetaSize = 4;
// offsets inside a Trajectory row (src/types.hpp:348-355)
constexpr int kTrajPos = 0;
constexpr int kTrajVel = 2 * kTrajLen;
constexpr int kTrajHead = 4 * kTrajLen;
constexpr int kTrajValid = 5 * kTrajLen;
constexpr int kTrajInv = 6 * kTrajLen;

// Parameters, src/init.hpp:111-127 (int32 flags instead of bool so the C ABI is unambiguous)
struct Params {
    float polylineReductionThreshold;
    float observationRadius;
    int32_t rewardType;
    float distanceToGoalThreshold;
    float distanceToExpertThreshold;
    int32_t collisionBehaviour;
    uint32_t maxNumControlledAgents;
    int32_t ignoreNonVehicles;
    int32_t roadObservationAlgorithm;
    int32_t initOnlyValidAgentsAtFirstStep;
    int32_t isStaticAgentControlled;
    int32_t enableLidar;
    int32_t disableClassicalObs;
    int32_t dynamicsModel;
    int32_t readFromTracksToPredict;
};

} // namespace gd
