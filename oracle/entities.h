// ORACLE — TEST INFRASTRUCTURE ONLY (see glsl_shim.h header). PARITY UNPINNED.
// Restatement of client/assets/shaders/compute_projectile.glsl and compute_entity.glsl.
#pragma once
#include "common.h"

namespace orc {

// set_data of compute_projectile.glsl:54-61 / compute_entity.glsl:53-60 is never called by either shader.

namespace compute_projectile {

struct Env {
    Images im;
    SimData sim_data;
    Projectile* projectiles;
    const Player* players;
    Collision* collisions;
    uint* collision_count;   // position taken: ONE global zero-initialised counter.
                             // The reference's `shared uint collision_count` is uninitialised and per-workgroup.
    uint collision_capacity;
    uint64_t* sample_counter;  // not in the reference: terrain samples visited (roofline model)
    uint64_t* los_iter_counter;
};

// compute_projectile.glsl:69-129
static inline bool los_check(const Env& e, vec3 start, vec3 end) {
    vec3 ray = normalize(end - start);
    float max_dist = length(end - start);

    uint offset = 0;
    if (ray.x < 0) offset += 1;
    if (ray.y < 0) offset += 2;
    if (ray.z < 0) offset += 4;
    float depth = 0;

    vec3 pos = start;

    for (uint i = 0; i < 20; i++) {
        if (e.los_iter_counter) (*e.los_iter_counter)++;
        vec3 floor_pos = floor(pos);
        uint voxel_data = get_data(e.im, e.sim_data, uvec3(floor_pos));
        vec3 v_min;
        vec3 v_max;
        uint voxel_material = voxel_data >> 24;
        if (voxel_material == MAT_AIR_OOB) {
            v_min = floor(pos / (float)CHUNK_SIZE) * (float)CHUNK_SIZE;
            v_max = v_min + vec3((float)CHUNK_SIZE);
        } else if (physics_properties[voxel_material].is_fluid) {
            uint dist = 0;
            if (physics_properties[voxel_material].is_data_standard_distance) {
                dist = get_dist(voxel_data, offset);
            }
            v_min = floor_pos - vec3((float)dist);
            v_max = floor_pos + vec3((float)(dist + 1));
        } else {
            return false;
        }
        vec3 delta = ray_box_dist(pos, ray, v_min, v_max);
        float dist_diff = min(delta.x, min(delta.y, delta.z));
        depth += dist_diff;
        if (depth > max_dist) {
            return true;
        }
        pos += ray * dist_diff;
        if (delta.x < delta.y && delta.x < delta.z) {
            if (ray.x > 0 && pos.x < v_max.x) {
                pos.x = v_max.x;
            } else if (ray.x < 0 && pos.x >= v_min.x) {
                pos.x = v_min.x - 0.001f;
            }
        } else if (delta.y < delta.z) {
            if (ray.y > 0 && pos.y < v_max.y) {
                pos.y = v_max.y;
            } else if (ray.y < 0 && pos.y >= v_min.y) {
                pos.y = v_min.y - 0.001f;
            }
        } else {
            if (ray.z > 0 && pos.z < v_max.z) {
                pos.z = v_max.z;
            } else if (ray.z < 0 && pos.z >= v_min.z) {
                pos.z = v_min.z - 0.001f;
            }
        }
    }

    return false;
}

// compute_projectile.glsl:133-243
static inline void main(const Env& e, uint projectile_idx) {
    if (projectile_idx >= e.sim_data.projectile_count) return;
    Projectile projectile = e.projectiles[projectile_idx];

    vec3 dir = quat_transform(projectile.dir, vec3(0.0f, 0.0f, 1.0f));
    vec3 right = quat_transform(projectile.dir, vec3(1.0f, 0.0f, 0.0f));
    vec3 up = quat_transform(projectile.dir, vec3(0.0f, 1.0f, 0.0f));

    vec3 projectile_vectors[] = {right, up, dir};

    vec3 adjusted_projectile_size = projectile.size.xyz();
    adjusted_projectile_size.z += e.sim_data.dt * projectile.vel / 2.0f;

    for (int i = 0; i < (int)e.sim_data.player_count; i++) {
        Player player = e.players[i];
        uint collision = 0;
        vec3 prev_collision = vec3(0.0f);
        for (int hitsphere_idx = 0; hitsphere_idx < HITSPHERE_COUNT; hitsphere_idx++) {
            Hitsphere hs = HITSPHERES(hitsphere_idx);
            Hitsphere transformed_hit_sphere = Hitsphere{player.pos.xyz() + hs.offset * player.size, hs.radius * player.size, hs.headshot};
            if (collision > 0) {
                if (length(projectile.pos.xyz() - transformed_hit_sphere.offset) > length(projectile.pos.xyz() - prev_collision)) {
                    continue;
                }
            }
            bool collide = true;
            for (int d = 0; d < 3; d++) {
                if (abs(dot(transformed_hit_sphere.offset, projectile_vectors[d]) - dot(projectile.pos.xyz(), projectile_vectors[d])) > adjusted_projectile_size[d] + transformed_hit_sphere.radius) {
                    collide = false;
                    break;
                }
            }
            if (!collide) {
                continue;
            }
            if (transformed_hit_sphere.headshot) {
                collision = 2;
            } else {
                collision = 1;
            }
            prev_collision = transformed_hit_sphere.offset;
        }
        if (collision > 0 && los_check(e, projectile.pos.xyz(), prev_collision)) {
            uint collision_idx;
#pragma omp atomic capture
            collision_idx = (*e.collision_count)++;
            if (collision_idx < e.collision_capacity)
                e.collisions[collision_idx] = Collision{projectile_idx, (uint)i, collision, dot(dir, prev_collision - projectile.pos.xyz())};
        }
    }

    if (projectile.should_collide_with_terrain != 1) {
        return;
    }

    vec4 gic = ceil(2.0f * projectile.size * sqrt(2.0f));
    ivec3 grid_iteration_count = ivec3(f2i(gic.x), f2i(gic.y), f2i(gic.z));
    grid_iteration_count.z = (f2i(ceil((2.0f * projectile.size.z + e.sim_data.dt * projectile.vel) * sqrt(2.0f))));
    vec3 grid_dist = 2.0f * projectile.size.xyz() / to_vec3(grid_iteration_count);
    grid_dist.z = (2.0f * projectile.size.z + e.sim_data.dt * projectile.vel) / (float)grid_iteration_count.z;
    vec3 start = projectile.pos.xyz() - dir * projectile.size.z - right * projectile.size.x - up * projectile.size.y;
    for (int i = 0; i <= grid_iteration_count.x; i++) {
        for (int j = 0; j <= grid_iteration_count.y; j++) {
            for (int k = 0; k <= grid_iteration_count.z; k++) {
                int mapped_i = grid_iteration_count.x / 2 + (2 * (i % 2) - 1) * ((i + 1) / 2);
                int mapped_j = grid_iteration_count.y / 2 + (2 * (j % 2) - 1) * ((j + 1) / 2);
                int mapped_k = grid_iteration_count.z / 2 + (2 * (k % 2) - 1) * ((k + 1) / 2);
                vec3 pos = start + dir * grid_dist.z * (float)mapped_k + right * grid_dist.x * (float)mapped_i + up * grid_dist.y * (float)mapped_j;
                ivec3 voxel_pos = ivec3(pos);
                if (e.sample_counter) (*e.sample_counter)++;
                uint data = get_data(e.im, e.sim_data, uvec3(voxel_pos));
                uint voxel_mat = data >> 24;
                if (physics_properties[voxel_mat].is_fluid) {
                    continue;
                } else if (voxel_mat == MAT_OOB) {
                    projectile.health = 0.0f;
                    projectile.chunk_update_pos.w = 0;
                    e.projectiles[projectile_idx] = projectile;
                    return;
                }

                float dist_past_bb = grid_dist.z * (float)mapped_k - 2.0f * projectile.size.z;
                vec3 delta = ray_box_dist(pos, -dir, to_vec3(voxel_pos), to_vec3(voxel_pos + ivec3(1)));
                float dist_diff = min(delta.x, min(delta.y, delta.z));
                if (dist_past_bb > 0.0f) {
                    projectile.pos += vec4(dir * (dist_past_bb - dist_diff), 0.0f);
                }

                if (projectile.wall_bounce == 1) {
                    vec3 new_dir = dir;
                    if (delta.x == dist_diff) {
                        new_dir.x *= -1.0f;
                    } else if (delta.y == dist_diff) {
                        new_dir.y *= -1.0f;
                    } else if (delta.z == dist_diff) {
                        new_dir.z *= -1.0f;
                    }
                    projectile.dir = quaternion_from_arc(vec3(0.0f, 0.0f, 1.0f), new_dir);
                    e.projectiles[projectile_idx] = projectile;
                    return;
                }

                if (projectile.damage > 0) {
                    // imageAtomicAdd(voxels, get_index(voxel_pos), int(projectile.damage))
                    uint64_t vox_index = get_index(e.im, e.sim_data, uvec3(voxel_pos));
                    uint add = (uint)f2i(projectile.damage);
#pragma omp atomic
                    e.im.voxels[vox_index] += add;
                }

                projectile.chunk_update_pos = uvec4((uint)voxel_pos.x, (uint)voxel_pos.y, (uint)voxel_pos.z, 1);

                projectile.health = 0.0f;
                e.projectiles[projectile_idx] = projectile;
                return;
            }
        }
    }
}

// dispatch(ceil(n/128)) of 128-wide workgroups (voxel_sim_manager.rs:944-950).
// Sequential: damage written by projectile p is visible to the terrain samples of later
// projectiles in the reference only by GPU scheduling accident; samples only test the
// material byte and damage adds never carry into it (max 24-bit headroom), so order is irrelevant
// for materials with is_data_damage. See tests for the distance-field caveat.
static inline void dispatch(const Env& e, bool parallel) {
#pragma omp parallel for schedule(dynamic, 256) if (parallel)
    for (int64_t p = 0; p < (int64_t)e.sim_data.projectile_count; p++) main(e, (uint)p);
}

}  // namespace compute_projectile

namespace compute_entity {

struct Env {
    Images im;
    SimData sim_data;
    Player* players;
};

// compute_entity.glsl:68-100
static inline bool can_step_up(const Env& e, Player player, uint component, vec3 player_move_pos) {
    vec3 start_pos =
        player.pos.xyz() + PLAYER_HITBOX_OFFSET() * player.size - 0.5f * player.size * PLAYER_HITBOX_SIZE();
    start_pos[component] = player_move_pos[component];
    start_pos[1] += 1.0f;
    int x_iter_count = f2i(floor(start_pos[(component + 1) % 3] + player.size * PLAYER_HITBOX_SIZE()[(component + 1) % 3]) - floor(start_pos[(component + 1) % 3]));
    int z_iter_count = f2i(floor(start_pos[(component + 2) % 3] + player.size * PLAYER_HITBOX_SIZE()[(component + 2) % 3]) - floor(start_pos[(component + 2) % 3]));

    vec3 x_vec = vec3(0.0f);
    vec3 z_vec = vec3(0.0f);
    x_vec[(component + 1) % 3] = 1.0f;
    z_vec[(component + 2) % 3] = 1.0f;
    for (int x_iter = 0; x_iter <= x_iter_count; x_iter++) {
        for (int z_iter = 0; z_iter <= z_iter_count; z_iter++) {
            vec3 pos = start_pos + float(x_iter) * x_vec + float(z_iter) * z_vec;
            uvec3 voxel_pos = uvec3(pos);
            uint voxel = get_data(e.im, e.sim_data, voxel_pos);
            uint voxel_material = voxel >> 24;
            if (!physics_properties[voxel_material].is_fluid) {
                return false;
            }
        }
    }
    return true;
}

// compute_entity.glsl:102-234
static inline void collide_player(const Env& e, Player& player, vec3 prev_collision_vec) {
    vec3 collision_corner_offset = sign(player.vel.xyz()) * PLAYER_HITBOX_SIZE() * 0.5f * player.size;
    vec3 distance_to_move = player.vel.xyz() * e.sim_data.dt;
    uint iteration_counter = 0;

    while (dot(distance_to_move, distance_to_move) > 0.0001f) {
        iteration_counter += 1;

        vec3 player_move_pos =
            player.pos.xyz() + PLAYER_HITBOX_OFFSET() * player.size + collision_corner_offset;
        vec3 vel_dir = normalize(distance_to_move);
        vec3 delta = ray_box_dist(player_move_pos, vel_dir, floor(player_move_pos), ceil(player_move_pos));
        if (isnan(delta.x)) {
            delta.x = 2.0f;
        }
        if (isnan(delta.y)) {
            delta.y = 2.0f;
        }
        if (isnan(delta.z)) {
            delta.z = 2.0f;
        }
        float dist_diff = min(delta.x, min(delta.y, delta.z));
        if (dist_diff == 0.0f) {
            dist_diff = length(distance_to_move);
            if (delta.x != 0.0f) {
                dist_diff = min(dist_diff, delta.x);
            }
            if (delta.y != 0.0f) {
                dist_diff = min(dist_diff, delta.y);
            }
            if (delta.z != 0.0f) {
                dist_diff = min(dist_diff, delta.z);
            }
        } else if (dist_diff > length(distance_to_move)) {
            player.pos.set_xyz(player.pos.xyz() + distance_to_move);
            break;
        }

        if (iteration_counter > 100) {
            break;
        }

        distance_to_move -= dist_diff * vel_dir;
        for (uint component = 0; component < 3; component++) {
            vec3 fake_pos = player.pos.xyz();
            fake_pos[component] += dist_diff * vel_dir[component];
            vec3 player_move_pos =
                fake_pos + PLAYER_HITBOX_OFFSET() * player.size + collision_corner_offset;
            if (delta[component] <= delta[(component + 1) % 3] && delta[component] <= delta[(component + 2) % 3]) {
                vec3 start_pos = fake_pos + PLAYER_HITBOX_OFFSET() * player.size - 0.5f * player.size * PLAYER_HITBOX_SIZE();
                start_pos[component] = player_move_pos[component];
                int x_iter_count = f2i(floor(start_pos[(component + 1) % 3] + player.size * PLAYER_HITBOX_SIZE()[(component + 1) % 3]) - floor(start_pos[(component + 1) % 3]));
                int z_iter_count = f2i(floor(start_pos[(component + 2) % 3] + player.size * PLAYER_HITBOX_SIZE()[(component + 2) % 3]) - floor(start_pos[(component + 2) % 3]));

                vec3 x_vec = vec3(0);
                vec3 z_vec = vec3(0);
                x_vec[(component + 1) % 3] = 1.0f;
                z_vec[(component + 2) % 3] = 1.0f;
                bool found_position = false;
                for (int x_iter = 0; x_iter <= x_iter_count; x_iter++) {
                    for (int z_iter = 0; z_iter <= z_iter_count; z_iter++) {
                        vec3 pos = start_pos + (float)x_iter * x_vec + (float)z_iter * z_vec;
                        uvec3 voxel_pos = uvec3(pos);
                        uint voxel = get_data(e.im, e.sim_data, voxel_pos);
                        uint voxel_material = voxel >> 24;
                        if (!physics_properties[voxel_material].is_fluid) {
                            if (component != 1 && prev_collision_vec[1] == 1 && (pos - start_pos).y < 1.0f && can_step_up(e, player, component, player_move_pos)) {
                                player.pos.set_xyz(fake_pos);
                                player.pos[1] += 1.0f;
                                found_position = true;
                                break;
                            }

                            player.vel[component] = 0.0f;
                            // apply friction
                            vec2 perp_vel = vec2(player.vel[(component + 1) % 3], player.vel[(component + 2) % 3]);
                            if (length(perp_vel) > 0.0f) {
                                float friction_factor = physics_properties[voxel_material].friction;
                                vec2 friction = vec2((friction_factor * 0.5f * normalize(perp_vel).x + friction_factor * perp_vel.x) * e.sim_data.dt,
                                                     (friction_factor * 0.5f * normalize(perp_vel).y + friction_factor * perp_vel.y) * e.sim_data.dt);
                                if (length(friction) > length(perp_vel)) {
                                    player.vel[(component + 1) % 3] = 0.0f;
                                    player.vel[(component + 2) % 3] = 0.0f;
                                } else {
                                    player.vel[(component + 1) % 3] -= friction.x;
                                    player.vel[(component + 2) % 3] -= friction.y;
                                }
                            }
                            player.collision_vec[component] = f2i(-sign(vel_dir[component]));
                            distance_to_move[component] = 0.0f;
                            found_position = true;
                            break;
                        }
                    }
                    if (found_position) {
                        break;
                    }
                }
                if (found_position) {
                    continue;
                }
            }
            player.pos.set_xyz(fake_pos);
        }
    }
}

// compute_entity.glsl:236-297
static inline void main(const Env& e, uint entity_idx) {
    if (entity_idx >= e.sim_data.player_count) return;
    Player player = e.players[entity_idx];

    // volume effects
    vec3 start_pos = player.pos.xyz() + player.size * PLAYER_HITBOX_OFFSET() - player.size * PLAYER_HITBOX_SIZE() / 2.0f;
    vec3 end_pos = player.pos.xyz() + player.size * PLAYER_HITBOX_OFFSET() + player.size * PLAYER_HITBOX_SIZE() / 2.0f;
    uvec3 start_voxel_pos = uvec3(start_pos);
    uvec3 iter_counts = uvec3(end_pos) - start_voxel_pos + uvec3(1);
    float nearby_density = 0.0f;
    vec3 directional_density = vec3(0.0f);
    for (int x = 0; x < (int)iter_counts.x; x++) {
        for (int y = 0; y < (int)iter_counts.y; y++) {
            for (int z = 0; z < (int)iter_counts.z; z++) {
                uvec3 voxel_pos = start_voxel_pos + uvec3((uint)x, (uint)y, (uint)z);
                vec3 overlapping_dimensions = min(end_pos, to_vec3(voxel_pos) + vec3(1.0f)) - max(start_pos, to_vec3(voxel_pos));
                float overlapping_volume = overlapping_dimensions.x * overlapping_dimensions.y * overlapping_dimensions.z;
                uint voxel = get_data(e.im, e.sim_data, voxel_pos);
                float density = physics_properties[voxel >> 24].density;
                nearby_density += overlapping_volume * density;
                directional_density += overlapping_volume * density * ((to_vec3(voxel_pos) + vec3(0.5f)) - (player.pos.xyz() + player.size * PLAYER_HITBOX_OFFSET())) / player.size;
            }
        }
    }
    nearby_density /= player.size * player.size * player.size * PLAYER_HITBOX_SIZE().x * PLAYER_HITBOX_SIZE().y * PLAYER_HITBOX_SIZE().z;
    directional_density = directional_density / (player.size * player.size * player.size * PLAYER_HITBOX_SIZE().x * PLAYER_HITBOX_SIZE().y * PLAYER_HITBOX_SIZE().z);

    if (player.has_world_collision == 1) {
        player.vel.set_xyz(player.vel.xyz() + (PLAYER_DENSITY - nearby_density) * player.gravity * 11.428571428571429f * e.sim_data.dt);
        if (length(directional_density) * e.sim_data.dt > 0.001f) {
            player.vel.set_xyz(player.vel.xyz() - 0.5f * directional_density * e.sim_data.dt);
        }
        if (dot(player.vel.xyz(), player.vel.xyz()) > 0.0f) {
            player.vel.set_xyz(player.vel.xyz() - (nearby_density * 0.0375f * player.vel.xyz() * length(player.vel.xyz()) * e.sim_data.dt + 0.2f * normalize(player.vel.xyz()) * e.sim_data.dt));
        }
        vec3 prev_collision_vec = to_vec3(player.collision_vec);
        player.collision_vec = ivec3(0);
        collide_player(e, player, prev_collision_vec);
    } else {
        player.pos += player.vel;
        player.vel *= 0.5f;
    }

    e.players[entity_idx] = player;
}

static inline void dispatch(const Env& e) {
    for (uint p = 0; p < e.sim_data.player_count; p++) main(e, p);
}

}  // namespace compute_entity
}  // namespace orc
