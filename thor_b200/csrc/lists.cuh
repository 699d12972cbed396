// device-side layouts shared by the list builder and the kernels
#pragma once
#include <stdint.h>
namespace thor {
// One <=16x16 luma tile of a prediction block, everything the kernel needs in one 32-byte record: picture HANDLES
// (the list-local indices are resolved at list creation), signed vectors (the sign flags already applied).
struct PredTile {
  uint16_t x, y;            // luma position of the tile
  uint8_t tw, th;           // luma size (4..16, multiples of 4)
  uint16_t dst, ref0, ref1; // picture handles; ref1 = 0xFFFF: uni-prediction
  int16_t mv0x, mv0y, mv1x, mv1y;
  uint32_t task;            // index of the task the tile belongs to
  uint32_t pad[2];
};
static_assert(sizeof(PredTile) == 32, "PredTile is read with two 16-byte loads");
}
