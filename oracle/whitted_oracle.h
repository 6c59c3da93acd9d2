/* whitted_oracle.h — TEST INFRASTRUCTURE. Entry points of the CPU oracle (see whitted_oracle.c).
 * It consumes the same flat views as the device path (include/b200rt.h) so that a parity test
 * feeds both from one ScenePack. */
#ifndef WHITTED_ORACLE_H
#define WHITTED_ORACLE_H

#include <stdint.h>
#include "../include/b200rt.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Renders one frame on `threads` CPU threads (tiles handed out as in WorkerThread.cpp:53-65).
 * Fills every non-NULL output; presented_rgb8 needs frame_buffer. config->tile_stride/phase are
 * honoured (pixels of other tiles are left untouched). Returns 0 or -1. */
int oracle_render(const b200rt_static_scene * scene, const b200rt_frame * frame, const b200rt_config * config,
                  const b200rt_outputs * outputs, int threads);

/* Gamma + FXAA pass (fragment_fxaa.glsl / fragment_identity.glsl) over a 0x00RRGGBB frame. */
void oracle_present(const uint32_t * frame_buffer, int width, int height, int enable_fxaa, uint8_t * rgb8);

/* pow(i / 255, 1 / 2.2) for i = 0..255: get_screen() of the shaders on 8-bit texels. */
void oracle_gamma_table(float table[256]);

#ifdef __cplusplus
}
#endif
#endif
