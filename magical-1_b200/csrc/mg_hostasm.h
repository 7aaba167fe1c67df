/* host-side assembly of stacked observations from a pinned mirror of the frame ring (mg_hostasm.cpp) */
#pragma once
#include <stdint.h>
struct MgHostAsm;
MgHostAsm *mg_hostasm_create(int n_threads);
void mg_hostasm_destroy(MgHostAsm *A);
/* one step: envs [0, n_envs) of obs_host are built from ring_host ([env][MG_MAX_RING][96*96*3]); group g of the stack is
 * the frame min(ages[g], since_reset[env]) steps old.  Workers start at once but touch env e only after a publish
 * with env_limit > e (its new frame has landed). */
void mg_hostasm_begin(MgHostAsm *A, const uint8_t *ring_host, uint8_t *obs_host, int n_envs, int newest_slot,
                      const int *ages, int n_groups, const uint8_t *since_reset);
void mg_hostasm_publish(MgHostAsm *A, int env_limit);
void mg_hostasm_finish(MgHostAsm *A);       /* the caller works too; returns when every env is built */
int mg_hostasm_threads(const MgHostAsm *A);
