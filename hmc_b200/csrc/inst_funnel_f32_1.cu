// Funnel, float32: compiled kernel shapes (EPL slots per lane, LPC lanes per chain).  Part 2 of 2: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(FunnelModel, HMCB_MODEL_FUNNEL, float, HMCB_F32, UNITM, EPL, LPC)
S(1, 4, 1) S(1, 20, 1) S(1, 50, 1) S(1, 28, 4) S(1, 4, 32) S(0, 10, 1)
