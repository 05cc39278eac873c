// Funnel, float64: compiled kernel shapes (EPL slots per lane, LPC lanes per chain).  Part 1 of 4: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(FunnelModel, HMCB_MODEL_FUNNEL, double, HMCB_F64, UNITM, EPL, LPC)
S(1, 2, 1) S(1, 52, 2) S(1, 14, 4) S(1, 8, 8) S(1, 4, 32)
