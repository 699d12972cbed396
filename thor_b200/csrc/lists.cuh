// device-side layouts shared by the list builder and the kernels
#pragma once
#include <stdint.h>
namespace thor {
struct PredTile { uint32_t task; uint8_t tx, ty, pad0, pad1; };   // tile offset in units of 16 luma samples
}
