/* ORACLE build support: the reference includes "plp_math.h" (conv.c:47) but uses nothing from it. */
