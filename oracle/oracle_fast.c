/*
 * oracle_fast.c -- the same restatement as oracle.c, compiled the way pgvector
 * is built (its Makefile: -O3-class optimisation with -ftree-vectorize
 * -fassociative-math -fno-signed-zeros -fno-trapping-math, recalled from the
 * public source; see oracle.c header).  Emits orc_topk_fast.  This is the
 * translation unit bench.py times as the CPU baseline; the parity checker is
 * the strict build of oracle.c.  TEST INFRASTRUCTURE ONLY.
 */
#define ORC_FAST 1
#include "oracle.c"
