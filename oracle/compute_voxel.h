// ORACLE — TEST INFRASTRUCTURE ONLY (see glsl_shim.h header). PARITY UNPINNED.
// Restatement of client/assets/shaders/compute_voxel.glsl (Margolus-block update),
// write_voxels.glsl, unload_chunks.glsl and complete_worldgen.glsl.
#pragma once
#include "common.h"

namespace orc {
namespace compute_voxel {

struct Invocation {
    const Images& im;
    const SimData& sim_data;
    ivec4* chunk_updates;  // binding 5; .w receives the modification flags
    uvec3 gl_WorkGroupID;
    uvec3 gl_LocalInvocationID;

    // compute_voxel.glsl:51-88
    void set_data(uvec3 global_pos, uint data) {
        uint og_voxel_data = get_data(im, sim_data, global_pos);
        if (og_voxel_data >> 24 == MAT_OOB) return;
        if (og_voxel_data >> 24 == MAT_AIR_OOB) return;
        if (og_voxel_data != data) {
            uvec3 pos_in_chunk = global_pos & 0xF;
            int modification_flags = 0x1;
            if (pos_in_chunk.x == 0 && gl_LocalInvocationID.x < 7) {
                modification_flags |= 0x2;
            }
            if (pos_in_chunk.y == 0 && gl_LocalInvocationID.y < 7) {
                modification_flags |= 0x4;
            }
            if (pos_in_chunk.z == 0 && gl_LocalInvocationID.z < 7) {
                modification_flags |= 0x8;
            }
            if (pos_in_chunk.x == 15) {
                modification_flags |= 0x10;
            }
            if (pos_in_chunk.x == 0 && gl_LocalInvocationID.x == 7) {
                modification_flags |= 0x10;
            }
            if (pos_in_chunk.y == 15) {
                modification_flags |= 0x20;
            }
            if (pos_in_chunk.y == 0 && gl_LocalInvocationID.y == 7) {
                modification_flags |= 0x20;
            }
            if (pos_in_chunk.z == 15) {
                modification_flags |= 0x40;
            }
            if (pos_in_chunk.z == 0 && gl_LocalInvocationID.z == 7) {
                modification_flags |= 0x40;
            }
            // atomicOr: invocations of one workgroup are run sequentially here
            chunk_updates[gl_WorkGroupID.x].w |= modification_flags;
        }
        im.voxels[get_index(im, sim_data, global_pos)] = data;
    }

    // compute_voxel.glsl:90-216
    void main() {
        const ivec3 gl_WorkGroupSize(8, 8, 8);
        ivec3 pos = 2 * gl_WorkGroupSize * chunk_updates[gl_WorkGroupID.x].xyz() + 2 * to_ivec3(gl_LocalInvocationID) + to_ivec3(sim_data.voxel_update_offset);
        uvec2 pos_data[2][2][2];
        for (uint i = 0; i < 2; i++) {
            for (uint j = 0; j < 2; j++) {
                for (uint k = 0; k < 2; k++) {
                    uint raw_voxel = get_data(im, sim_data, uvec3(pos + ivec3(i, j, k)));
                    pos_data[i][j][k] = uvec2(raw_voxel >> 24, raw_voxel & 0xFFFFFF);
                }
            }
        }

        for (uint i = 0; i < 2; i++) {
            for (uint j = 0; j < 2; j++) {
                for (uint k = 0; k < 2; k++) {
                    if (physics_properties[pos_data[i][j][k].x].is_data_damage) {
                        if (pos_data[i][j][k].y >= material_damage_threshhold[pos_data[i][j][k].x]) {
                            pos_data[i][j][k] = uvec2(MAT_AIR, 0);
                        }
                    }
                }
            }
        }

        for (uint i = 0; i < 2; i++) {
            for (uint k = 0; k < 2; k++) {
                uint bottom_voxel = pos_data[i][0][k].x;
                uint top_voxel = pos_data[i][1][k].x;
                if (bottom_voxel == MAT_AIR && top_voxel == MAT_WATER) {
                    pos_data[i][0][k] = uvec2(top_voxel, 0);
                    pos_data[i][1][k] = uvec2(bottom_voxel, 0);
                }
            }
        }

        for (uint i = 0; i < 2; i++) {
            for (uint k = 0; k < 2; k++) {
                uint bottom_voxel = pos_data[1 - i][0][k].x;
                uint top_voxel = pos_data[i][1][k].x;
                if (bottom_voxel == MAT_AIR && top_voxel == MAT_WATER) {
                    pos_data[1 - i][0][k] = uvec2(top_voxel, 0);
                    pos_data[i][1][k] = uvec2(bottom_voxel, 0);
                }
            }
        }

        for (uint i = 0; i < 2; i++) {
            for (uint k = 0; k < 2; k++) {
                uint bottom_voxel = pos_data[i][0][1 - k].x;
                uint top_voxel = pos_data[i][1][k].x;
                if (bottom_voxel == MAT_AIR && top_voxel == MAT_WATER) {
                    pos_data[i][0][1 - k] = uvec2(top_voxel, 0);
                    pos_data[i][1][k] = uvec2(bottom_voxel, 0);
                }
            }
        }

        for (uint i = 0; i < 2; i++) {
            uint to_voxel = pos_data[pos.y % 2][1][i].x;
            uint from_voxel = pos_data[(1 + pos.y) % 2][1][i].x;
            if (from_voxel == MAT_WATER && to_voxel == MAT_AIR) {
                pos_data[pos.y % 2][1][i] = uvec2(from_voxel, 0);
                pos_data[(1 + pos.y) % 2][1][i] = uvec2(to_voxel, 0);
            }
        }

        for (uint i = 0; i < 2; i++) {
            uint to_voxel = pos_data[i][1][pos.y % 2].x;
            uint from_voxel = pos_data[i][1][(1 + pos.y) % 2].x;
            if (from_voxel == MAT_WATER && to_voxel == MAT_AIR) {
                pos_data[i][1][pos.y % 2] = uvec2(from_voxel, 0);
                pos_data[i][1][(1 + pos.y) % 2] = uvec2(to_voxel, 0);
            }
        }

        for (uint i = 0; i < 2; i++) {
            for (uint j = 0; j < 2; j++) {
                for (uint k = 0; k < 2; k++) {
                    if (pos_data[i][j][k].x == MAT_AIR) {
                        int offset = 3 * int(((k << 2) | (j << 1) | i));
                        ivec3 d = ivec3(1) - ivec3(i, j, k) * ivec3(2);
                        uint direction_dist = 7;
                        for (uint m = 1; m < 8; m++) {
                            ivec3 dir = d * ivec3(m % 2, (m / 2) % 2, m / 4) + ivec3(i, j, k);
                            if (pos_data[dir.x][dir.y][dir.z].x == MAT_OOB) {
                                direction_dist = min(direction_dist, bitfieldExtract(pos_data[i][j][k].y, offset, 3));
                                continue;
                            } else if (pos_data[dir.x][dir.y][dir.z].x == MAT_AIR) {
                                direction_dist = min(direction_dist, bitfieldExtract(pos_data[dir.x][dir.y][dir.z].y, offset, 3) + 1);
                                continue;
                            } else if (pos_data[dir.x][dir.y][dir.z].x == MAT_AIR_OOB) {
                                continue;
                            }
                            direction_dist = 0;
                            break;
                        }
                        pos_data[i][j][k].y = bitfieldInsert(pos_data[i][j][k].y, direction_dist, offset, 3);
                    } else if (physics_properties[pos_data[i][j][k].x].is_data_standard_distance) {
                        int offset = 3 * int(((k << 2) | (j << 1) | i));
                        ivec3 d = ivec3(1) - ivec3(i, j, k) * ivec3(2);
                        uint direction_dist = 7;
                        for (uint m = 1; m < 8; m++) {
                            ivec3 dir = d * ivec3(m % 2, (m / 2) % 2, m / 4) + ivec3(i, j, k);
                            if (pos_data[dir.x][dir.y][dir.z].x == MAT_OOB) {
                                direction_dist = min(direction_dist, bitfieldExtract(pos_data[i][j][k].y, offset, 3));
                                continue;
                            } else if (pos_data[dir.x][dir.y][dir.z].x == pos_data[i][j][k].x) {
                                direction_dist = min(direction_dist, bitfieldExtract(pos_data[dir.x][dir.y][dir.z].y, offset, 3) + 1);
                                continue;
                            }
                            direction_dist = 0;
                            break;
                        }
                        pos_data[i][j][k].y = bitfieldInsert(pos_data[i][j][k].y, direction_dist, offset, 3);
                    }
                }
            }
        }

        for (uint i = 0; i < 2; i++) {
            for (uint j = 0; j < 2; j++) {
                for (uint k = 0; k < 2; k++) {
                    set_data(uvec3(pos + ivec3(i, j, k)), (pos_data[i][j][k].x << 24) | pos_data[i][j][k].y);
                }
            }
        }
    }
};

// vkCmdDispatch(chunk_update_count, 1, 1) with local size 8x8x8 (voxel_sim_manager.rs:994-1000).
// Workgroups of one phase touch disjoint voxels, so they may run in parallel.
static inline void dispatch(const Images& im, const SimData& sim_data, ivec4* chunk_updates, uint count) {
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t wg = 0; wg < (int64_t)count; wg++) {
        for (uint lx = 0; lx < 8; lx++)
            for (uint ly = 0; ly < 8; ly++)
                for (uint lz = 0; lz < 8; lz++) {
                    Invocation inv{im, sim_data, chunk_updates, uvec3((uint)wg, 0, 0), uvec3(lx, ly, lz)};
                    inv.main();
                }
    }
}

}  // namespace compute_voxel

namespace write_voxels {
// write_voxels.glsl:40-45 — one invocation per queued write, no bounds check, no table check.
static inline void dispatch(const Images& im, const SimData& sim_data, const uvec4* voxel_writes) {
    for (uint gid = 0; gid < sim_data.voxel_write_count; gid++) {
        im.voxels[get_index(im, sim_data, voxel_writes[gid].xyz())] = voxel_writes[gid].w;
    }
}
}  // namespace write_voxels

namespace unload_chunks {
// unload_chunks.glsl:19-26; grid = rs[(c+1)%3]/16 x rs[(c+2)%3]/16 workgroups of 16x16
// (voxel_sim_manager.rs:959-971) — i.e. only floor(rs/16)*16 rows are cleared, as in the reference.
static inline void dispatch(const Images& im, const SimData& sim_data) {
    uint c = sim_data.unload_component;
    uint nx = (sim_data.render_size[(c + 1) % 3] / 16) * 16;
    uint ny = (sim_data.render_size[(c + 2) % 3] / 16) * 16;
    for (uint gx = 0; gx < nx; gx++)
        for (uint gy = 0; gy < ny; gy++) {
            ivec3 coords = ivec3(0);
            coords[c] = sim_data.unload_index % int(sim_data.render_size[c]);
            coords[(c + 1) % 3] = int(gx);
            coords[(c + 2) % 3] = int(gy);
            im.chunks[table_index(uvec3(coords), sim_data.render_size)] = 0;
        }
}
}  // namespace unload_chunks

namespace complete_worldgen {
// complete_worldgen.glsl:29-36
static inline void dispatch(const Images& im, const SimData& sim_data, const ivec4* chunk_loads, const uint* load_results) {
    for (uint gid = 0; gid < sim_data.worldgen_count; gid++) {
        uvec3 pos = 8u * uvec3(chunk_loads[8 * gid].xyz());
        uvec4 indicies = get_indicies(pos, sim_data.render_size);
        im.chunks[table_index(indicies.xyz(), sim_data.render_size)] = load_results[gid];
    }
}
}  // namespace complete_worldgen

}  // namespace orc
