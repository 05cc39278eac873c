// Funnel, float32: compiled kernel shapes (EPL slots per lane, LPC lanes per chain).  Part 1 of 2: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(FunnelModel, HMCB_MODEL_FUNNEL, float, HMCB_F32, UNITM, EPL, LPC)
S(1, 10, 1) S(1, 26, 2) S(1, 52, 2) S(1, 8, 32) S(0, 14, 4) S(0, 8, 32)
