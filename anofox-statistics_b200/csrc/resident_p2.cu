#include "resident.cuh"
#include "resident_launch.h"
namespace b200 {
B200_DEFINE_RESIDENT(2)
}
