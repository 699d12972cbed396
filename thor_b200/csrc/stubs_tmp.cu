#include "internal.cuh"
extern "C" {
size_t thorcuda_sad_table_bytes(int) { return 0; }
void *thorcuda_sad_table_devptr(int) { return nullptr; }
int thorcuda_sad_table_fetch(int, void *, size_t) { return thor::fail("not built yet"); }
int thorcuda_encode_blocks(thorcuda_pic, thorcuda_pic, thorcuda_pic, const thorcuda_tb_task *, int, size_t, int16_t *, uint8_t *, uint32_t *) { return thor::fail("not built yet"); }
int thorcuda_reconstruct_blocks(thorcuda_pic, thorcuda_pic, const thorcuda_tb_task *, int, const int16_t *, size_t) { return thor::fail("not built yet"); }
}
