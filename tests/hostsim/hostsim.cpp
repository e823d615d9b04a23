// hostsim.cpp -- TEST SCAFFOLDING, not product: the per-env arithmetic of legged_tracking_b200/csrc/lt_core.cuh
// compiled for the host and driven by a plain sequential loop, so that the scalar logic the CUDA kernels share
// (ordering of reset / reward / observation semantics, rounding of each op) can be checked against the golden
// fixtures in the GPU-less build container.  The kernels' own choreography (tile staging, coalesced row copies,
// heightfield gathers, compaction) is only exercised by the `-m gpu` tests.  Nothing under legged_tracking_b200/
// loads this library.
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../legged_tracking_b200/csrc/lt_core.cuh"

extern "C" {

int lths_config_bytes(void) { return (int)sizeof(lt_config); }

int lths_torques(const lt_config* cfg, const lt_tables* tb, const lt_state_ptrs* stp, const float* dof_state,
                 const float* actions_raw, int substep, int32_t* counters) {
    const lt_config& c = *cfg;
    const lt_state_ptrs& st = *stp;
    lt_act_weights W;
    memcpy(W.w0, tb->act_w0, sizeof W.w0); memcpy(W.b0, tb->act_b0, sizeof W.b0);
    memcpy(W.w1, tb->act_w1, sizeof W.w1); memcpy(W.b1, tb->act_b1, sizeof W.b1);
    memcpy(W.w2, tb->act_w2, sizeof W.w2); memcpy(W.b2, tb->act_b2, sizeof W.b2);
    const size_t n = (size_t)c.num_envs * c.num_dof;
    const int Lg = c.lag_timesteps + 1;
    const int head = (counters[2] + substep) % Lg, rd = (head + 1) % Lg;
    for (size_t s = 0; s < n; ++s) {
        const int dof = (int)(s % c.num_dof);
        lt_torque_io io;
        io.action_raw = actions_raw[s];
        io.dof_pos = dof_state[2 * s];
        io.dof_vel = dof_state[2 * s + 1];
        io.action_scaled = lt_scaled_action(c, io.action_raw, dof, io.action_clipped);
        io.lag_out = 0.0f;
        if (c.randomize_lag_timesteps) {
            st.lag_buffer[(size_t)head * n + s] = io.action_scaled;
            io.lag_out = st.lag_buffer[(size_t)rd * n + s];
        }
        io.motor_offset = st.motor_offsets[s];
        io.motor_strength = st.motor_strengths[s];
        io.kp_factor = st.Kp_factors[s];
        io.kd_factor = st.Kd_factors[s];
        io.err_last = st.joint_pos_err_last[s];
        io.err_last_last = st.joint_pos_err_last_last[s];
        io.vel_last = st.joint_vel_last[s];
        io.vel_last_last = st.joint_vel_last_last[s];
        lt_torque_sample(c, W, dof, io);
        if (c.control_type == LT_CONTROL_ACTUATOR_NET) {
            st.joint_pos_err_last[s] = io.err_last;
            st.joint_pos_err_last_last[s] = io.err_last_last;
            st.joint_vel_last[s] = io.vel_last;
            st.joint_vel_last_last[s] = io.vel_last_last;
        }
        if (substep == 0) st.actions[s] = io.action_clipped;
        st.joint_pos_target[s] = io.joint_pos_target;
        st.torques[s] = io.torque;
    }
    return 0;
}

int lths_post_physics(const lt_config* cfg, const lt_tables* tb, const lt_state_ptrs* stp, const lt_sim_ptrs* simp,
                      const lt_terrain_ptrs* terp, const lt_step_args* argsp, int32_t* counters, int n_substeps) {
    const lt_config& c = *cfg;
    const lt_state_ptrs& st = *stp;
    const lt_sim_ptrs& sim = *simp;
    const lt_terrain_ptrs& ter = *terp;
    lt_step_args args = *argsp;
    lt_layout L;
    lt_make_layout(c, L);
    const int N = c.num_envs, nd = c.num_dof, no = c.num_obs, H = c.obs_history_length;
    const int A13 = c.num_actors * 13, B3 = c.num_bodies * 3;
    lt_step_scalars sc;
    memset(&sc, 0, sizeof sc);
    sc.step = (uint32_t)counters[0];
    if (args.new_gravity) { sc.set_gravity = 1; for (int k = 0; k < 3; ++k) sc.new_gravity[k] = args.new_gravity[k]; }
    std::vector<float> term_sums(LT_MAX_TERMS, 0.0f);
    sc.term_sums = term_sums.data();
    // sign pre-pass (lt_sign_kernel)
    for (int t = 0; t < c.n_terms; ++t) {
        if (c.term_bucket[t] != LT_BUCKET_BY_SUM) continue;
        double acc = 0.0;
        for (int i = 0; i < N; ++i) {
            const float* root = sim.root_states + (size_t)i * A13;
            lt_pose P;
            lt_derive_pose(st, c, i, root, st.curr_pose_index[i], P);
            lt_reward_ctx rc;
            rc.dof = sim.dof_state + (size_t)i * nd * 2; rc.last_dof_vel = st.last_dof_vel + (size_t)i * nd;
            rc.torques = st.torques + (size_t)i * nd; rc.actions = st.actions + (size_t)i * nd;
            rc.last_actions = st.last_actions + (size_t)i * nd; rc.cinfo = lt_contact_summary(c, sim.contact_forces + (size_t)i * B3);
            rc.root_z = root[2];
            for (int k = 0; k < 3; ++k) { rc.blv[k] = P.blv[k]; rc.bav[k] = P.bav[k]; rc.pg[k] = P.pg[k];
                                          rc.rel_lin[k] = P.rel_lin[k]; rc.rel_rot[k] = P.rel_rot[k]; rc.loc_lin[k] = P.rel_lin[k]; }
            rc.ep_f = (float)(st.episode_length_buf[i] + 1); rc.reached = 0; rc.n_collisions = 0;
            acc += lt_reward_term(c, st, i, c.term_ids[t], rc, lt_ext_row(c, args.extra_reward, t), true) * c.term_scales[t];
        }
        term_sums[t] = (float)acc;
    }
    std::vector<uint8_t> ix(L.nxf + 1), iy(c.n_points_y + 1);
    std::vector<float> la_pre(nd);
    for (int i = 0; i < N; ++i) {
        lt_env_rows R;
        R.root = sim.root_states + (size_t)i * A13;
        R.arrow = c.num_actors > 1 ? R.root + (c.num_actors - 1) * 13 : nullptr;
        R.dof = sim.dof_state + (size_t)i * nd * 2;
        R.cinfo = lt_contact_summary(c, sim.contact_forces + (size_t)i * B3);
        R.actions = st.actions + (size_t)i * nd;
        for (int d = 0; d < nd; ++d) la_pre[d] = st.last_actions[(size_t)i * nd + d];
        R.last_actions = la_pre.data();
        R.last_dof_vel = st.last_dof_vel + (size_t)i * nd;
        R.torques = st.torques + (size_t)i * nd;
        R.obs_row = st.obs_buf + (size_t)i * no; R.ix = ix.data(); R.iy = iy.data();
        R.es = st.episode_sums + i; R.cs = st.command_sums ? st.command_sums + i : nullptr; R.sum_stride = (size_t)N;
        const lt_env_result r = lt_env_step(c, L, st, ter, args, sc, i, R);
        // history shifts (kernel phase 3)
        for (int d = 0; d < nd; ++d) {
            const size_t k = (size_t)i * nd + d;
            st.last_last_actions[k] = r.reset ? 0.0f : la_pre[d];
            st.last_actions[k] = st.actions[k];
            st.last_last_joint_pos_target[k] = st.last_joint_pos_target[k];
            st.last_joint_pos_target[k] = st.joint_pos_target[k];
            st.last_dof_vel[k] = R.dof[2 * d + 1];
        }
        for (int j = 0; j < 6; ++j) st.last_root_vel[(size_t)i * 6 + j] = R.root[7 + j];
        // observation row (kernel phase 4)
        float* obs_row = st.obs_buf + (size_t)i * no;
        float* hist_row = st.obs_history + (size_t)i * no * H;
        if (H > 1) memmove(hist_row, hist_row + no, sizeof(float) * (size_t)(H - 1) * no);
        float* hist_new = hist_row + (size_t)(H - 1) * no;
        const float* tile_base = (c.plane || !c.observe_heights) ? nullptr
                                 : ter.tile_heights + (size_t)ter.env_tile[i] * 2 * L.tile_cells;
        for (int col = 0; col < no; ++col) {
            float v;
            const int hc = col - L.h0;
            if (hc >= 0 && hc < L.n_heights) {
                const int per_layer = L.nxf * c.n_points_y;
                const int layer = hc / per_layer, rem = hc % per_layer, a = rem / c.n_points_y, b = rem % c.n_points_y;
                float h;
                if (c.plane) h = layer == 0 ? 1.0f : 0.0f;
                else h = tile_base[(size_t)layer * L.tile_cells + (int)ix[a] * c.tile_cols + (int)iy[b]];
                v = lt_height_obs(c, h, r.z_post, r.cam_z);
            } else {
                v = obs_row[col];
            }
            float u = 0.5f;
            if (c.add_noise) {
                if (args.noise_uniforms) u = args.noise_uniforms[(size_t)i * no + col];
                else u = lt_noise_uniform(c, L, (uint32_t)i, col, sc.step);
            }
            v = lt_obs_finish(c, v, u, c.add_noise ? tb->noise_scale_vec[col] : 0.0f);
            obs_row[col] = v;
            hist_new[col] = v;
        }
    }
    // lt_compact_kernel
    int total = 0;
    for (int i = 0; i < N; ++i) if (st.reset_buf[i]) st.reset_env_ids[total++] = i;
    st.reset_count[0] = total;
    if (c.send_timeouts && total > 0) memcpy(st.time_outs_latched, st.time_out_buf, (size_t)N);
    counters[0] += 1;
    counters[1] += n_substeps;
    counters[2] = (counters[2] + n_substeps) % L.lag_slots;
    return 0;
}

}  // extern "C"
