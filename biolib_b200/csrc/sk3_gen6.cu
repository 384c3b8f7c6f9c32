// general form of the sk3 kernel, blocks of 6 m-mer starts per lane (see sk3_gen.cuh, sk3_device.cuh)
#include "sk3_gen.cuh"
BLG_SK3_GEN_DEFINE(6)
