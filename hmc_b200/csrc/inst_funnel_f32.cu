// Funnel, float32: compiled kernel shapes (EPL slots per lane, LPC lanes per chain).
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(FunnelModel, HMCB_MODEL_FUNNEL, float, HMCB_F32, UNITM, EPL, LPC)
S(1, 4, 1) S(1, 10, 1) S(1, 20, 1) S(1, 50, 1) S(1, 26, 2) S(1, 52, 2) S(1, 28, 4) S(1, 4, 32) S(1, 8, 32)
S(0, 10, 1) S(0, 14, 4) S(0, 8, 32)
