#include "fused_launch.h"
namespace b200 {
B200_DEFINE_SMALL(2)
}
