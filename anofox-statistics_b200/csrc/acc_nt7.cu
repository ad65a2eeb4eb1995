#include "acc_launch.h"
namespace b200 {
B200_DEFINE_NT(7)
}
