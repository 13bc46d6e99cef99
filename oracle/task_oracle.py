"""Single-environment CPU oracle of the task layer. TEST INFRASTRUCTURE ONLY.

Restates, in plain numpy around `OraclePhysics`, what the reference's Python does per env step:
composer Environment.step hook order, `Piano._update_key_state`
(reference piano.py:178-192), `PianoWithShadowHands.before_step/after_step/_compute_*_reward/
_update_goal_state/_update_fingering_state` (piano_with_shadow_hands.py:176-412),
`SelfActuatedPiano` (self_actuated_piano.py:143-228), `CompositeReward.compute`
(composite_reward.py:46-56) and `dm_control.utils.rewards.tolerance` (gaussian sigmoid,
value_at_margin 0.1; third-party, restated from its published formula).
"parity unpinned" vs the real reference for everything that depends on hand physics.
"""

from __future__ import annotations

from collections import OrderedDict

import numpy as np
from scipy.optimize import linear_sum_assignment

from .oracle_physics import OraclePhysics

_KEY_THRESHOLD = 0.00872665
_SUSTAIN_THRESHOLD = 0.5
_FINGER_CLOSE_ENOUGH_TO_KEY = 0.01
_KEY_CLOSE_ENOUGH_TO_PRESSED = 0.05
_EPS = 1e-6


def tolerance(x, bounds=(0.0, 0.0), margin=0.0, value_at_margin=0.1):
    """dm_control.utils.rewards.tolerance with sigmoid='gaussian'."""
    lower, upper = bounds
    x = np.asarray(x, np.float64)
    in_bounds = np.logical_and(lower <= x, x <= upper)
    if margin == 0:
        value = np.where(in_bounds, 1.0, 0.0)
    else:
        d = np.where(x < lower, lower - x, x - upper) / margin
        scale = np.sqrt(-2 * np.log(value_at_margin))
        value = np.where(in_bounds, 1.0, np.exp(-0.5 * (d * scale) ** 2))
    return float(value) if np.isscalar(x) or value.ndim == 0 else value


class _Base:
    def __init__(self, task):
        self.task = task
        self.model = task.model
        self.physics = OraclePhysics(task.model)
        self.notes, self.sustains = task._notes, task._sustains
        self.L = task._n_steps_lookahead
        self.nsub = task.physics_steps_per_control_step
        self.range_hi = self.model.a["jnt_range"][:88, 1].copy()

    def _goal_obs(self):
        if self.t_idx == len(self.notes):
            return
        g = np.zeros((self.L + 1, 89))
        for i, t in enumerate(range(self.t_idx, min(self.t_idx + self.L + 1, len(self.notes)))):
            g[i, [n.key for n in self.notes[t]]] = 1.0
            g[i, -1] = self.sustains[t]
        self.goal_state = g


class SelfActuatedPianoOracle(_Base):
    def __init__(self, task):
        super().__init__(task)
        self.ctrl_mid = 0.5 * self.model.a["act_ctrlrange"][:, 1]
        self.reset()

    def reset(self):
        self.physics.reset()
        self.t_idx = 0
        self.activation = np.zeros(88, bool)
        self.sustain_state = 0.0
        self.sustain_activation = False
        self._goal_obs()
        return self.observation()

    def observation(self):
        return OrderedDict([("piano/activation", self.activation.astype(np.float64)),
                            ("piano/sustain_activation", np.array([float(self.sustain_activation)])),
                            ("goal", self.goal_state.ravel().copy())])

    def step(self, action):
        action = np.asarray(action, np.float64)
        p = self.physics
        p.ctrl[:] = action[:-1]
        self.sustain_state = action[-1]
        for _ in range(self.nsub):
            p.step()
            self.activation = p.ctrl >= self.ctrl_mid
            self.sustain_activation = self.sustain_state >= _SUSTAIN_THRESHOLD
        self.t_idx += 1
        done = (self.t_idx - 1) == len(self.notes) - 1
        goal_current = self.goal_state[0]
        self._goal_obs()
        pred = np.concatenate([self.activation, [self.sustain_activation]]).astype(np.float64)
        if self.task.reward_type.value == "negative_l2":
            reward = -np.sqrt(np.sum((pred - goal_current) ** 2))
        else:
            reward = np.sum(goal_current * np.log(pred + _EPS) + (1 - goal_current) * np.log(1 - pred + _EPS))
        return self.observation(), float(reward), bool(done)


class PianoWithShadowHandsOracle(_Base):
    def __init__(self, task, legacy_step: bool = True):
        super().__init__(task)
        # dm_control Physics.step: legacy_step=True runs mj_step2 ... mj_step1 (derived quantities at the integrated
        # state), False runs plain mj_step (they stay those of the last mj_forward, i.e. of the last substep's start)
        self.legacy_step = bool(legacy_step)
        a = self.model.a
        self.has_fingering = not task._disable_fingering_reward
        # one-hand task (piano_with_one_shadow_hand.py): "left" / "right", None for the two-hands task
        self.side = getattr(task, "_side", None)
        self.act_lo, self.act_hi = a["act_ctrlrange"][:, 0].copy(), a["act_ctrlrange"][:, 1].copy()
        # hand.joints order: 24 menagerie joints, then forearm_tx, forearm_ty
        order = list(range(2, 26)) + [0, 1]
        self.rh_joints = 88 + np.array(order)
        self.lh_joints = 114 + np.array(order)
        self.key_geom_pos_offset = np.stack([0.35 * a["geom_size"][:88, 0], np.zeros(88),
                                             0.5 * a["geom_size"][:88, 2]], 1)
        cls = a["geom_cls"]
        self.rh_forearm = set(np.flatnonzero(cls == 4))
        self.lh_forearm = set(np.flatnonzero(cls == 5))
        self.reset()

    def reset(self, hand_offset: float = 0.0):
        """hand_offset: the y shift _randomize_initial_hand_positions would draw for this episode
        (piano_with_shadow_hands.py:491-499); the caller supplies the draw."""
        p = self.physics
        p.shift_hands(hand_offset)
        p.reset()
        p.forward_position(True)
        self.t_idx = 0
        self.discount = 1.0
        self.sustain_state = 0.0
        self._key_state()
        self._goal_obs()
        self._fingering_obs()
        self.reward_terms = OrderedDict()
        return self.observation()

    def _key_state(self):
        q = self.physics.qpos[:88]
        self.state = np.clip(q, 0.0, self.range_hi)
        self.normalized_state = self.state / self.range_hi
        self.activation = np.abs(self.state - self.range_hi) <= _KEY_THRESHOLD
        self.sustain_activation = self.sustain_state >= _SUSTAIN_THRESHOLD

    def _fingering_obs(self):
        if self.t_idx == len(self.notes):
            return
        self.rh_keys, self.lh_keys = [], []
        for note in self.notes[self.t_idx]:
            if note.fingering < 5:
                self.rh_keys.append((note.key, note.fingering))
            else:
                self.lh_keys.append((note.key, note.fingering - 5))
        if self.side is not None:  # piano_with_one_shadow_hand.py:297-318: this hand's notes only, a 5-vector
            if self.side == "right":
                self.lh_keys = []
            else:
                self.rh_keys = []
            f = np.zeros(5)
            for _, finger in self.rh_keys + self.lh_keys:
                f[finger] = 1.0
            self.fingering_state = f
            return
        f = np.zeros((2, 5))
        for hand, keys in enumerate([self.rh_keys, self.lh_keys]):
            for _, finger in keys:
                f[hand, finger] = 1.0
        self.fingering_state = f

    def observation(self):
        p = self.physics
        if self.side is not None:  # :320-361: the one hand's joints_pos and position (xpos of its root body)
            pre = "lh" if self.side == "left" else "rh"
            body = self.model.names["body"].index(f"{pre}_shadow_hand/{pre}_forearm")
            obs = OrderedDict([("piano/state", self.normalized_state.copy()),
                               ("piano/sustain_state", np.array([self.sustain_state])),
                               (f"{pre}_shadow_hand/joints_pos",
                                p.qpos[self.lh_joints if self.side == "left" else self.rh_joints].copy()),
                               (f"{pre}_shadow_hand/position", p.xpos[body].copy()),
                               ("goal", self.goal_state.ravel().copy())])
            if self.has_fingering:
                obs["fingering"] = self.fingering_state.ravel().copy()
            return obs
        obs = OrderedDict([("piano/state", self.normalized_state.copy()),
                           ("piano/sustain_state", np.array([self.sustain_state])),
                           ("rh_shadow_hand/joints_pos", p.qpos[self.rh_joints].copy()),
                           ("lh_shadow_hand/joints_pos", p.qpos[self.lh_joints].copy()),
                           ("goal", self.goal_state.ravel().copy())])
        if self.has_fingering:
            obs["fingering"] = self.fingering_state.ravel().copy()
        return obs

    def canonical_to_spec(self, action):
        lo = np.concatenate([self.act_lo, [0.0]])
        hi = np.concatenate([self.act_hi, [1.0]])
        return 0.5 * (np.asarray(action, np.float64) + 1.0) * (hi - lo) + lo

    def step(self, action):
        action = np.asarray(action, np.float64)
        p = self.physics
        p.ctrl[:] = action[:-1]
        self.sustain_state = action[-1]
        a = self.model.a
        j0, j1 = a["act_trn0"], a["act_trn1"]
        for i in range(self.nsub):
            if i == self.nsub - 1:  # what the last mj_step's forward pass sees
                q_start, v_start = p.qpos.copy(), p.qvel.copy()
            p.step()
            self._key_state()
        force = p.actuator_force.copy()  # from the last mj_step2 (pre-integration)
        if self.legacy_step:
            p.forward_position(True)  # trailing mj_step1: positions + contacts at the new state
            vel = p.actuator_velocity.copy()
        else:
            # no trailing mj_step1: xpos / site_xpos / geom_xpos / contacts / actuator_velocity are the last mj_step's
            q_new = p.qpos.copy()
            p.qpos[:] = q_start
            p.forward_position(True)
            p.qpos[:] = q_new
            vel = v_start[j0] + np.where(j1 >= 0, v_start[np.maximum(j1, 0)], 0.0)
        return self._after_physics(force, vel)

    def epilogue_at(self, qpos, qvel, actuator_force, sustain_state):
        """The task layer alone on a GIVEN end-of-step state (tests: the CUDA step's own qpos / qvel and
        actuatorfrc sensor in, every reward term / observation / flag out - no physics drift in between).
        The oracle's frame counter and cached goal / fingering must be those of the step being closed."""
        p = self.physics
        p.qpos[:] = np.asarray(qpos, np.float64)
        p.qvel[:] = np.asarray(qvel, np.float64)
        self.sustain_state = float(sustain_state)
        self._key_state()
        p.forward_position(True)
        a = self.model.a
        j0, j1 = a["act_trn0"], a["act_trn1"]
        vel = p.qvel[j0] + np.where(j1 >= 0, p.qvel[np.maximum(j1, 0)], 0.0)
        return self._after_physics(np.asarray(actuator_force, np.float64), vel)

    def _after_physics(self, force, vel):
        p = self.physics
        # after_step
        self.t_idx += 1
        should_terminate = (self.t_idx - 1) == len(self.notes) - 1
        goal_current = self.goal_state[0]
        rh_cur, lh_cur = self.rh_keys, self.lh_keys
        failure = bool(self.activation[np.flatnonzero(1 - goal_current[:-1])].any())
        self._goal_obs()
        self._fingering_obs()
        # rewards, in insertion order
        terms = OrderedDict()
        on = np.flatnonzero(goal_current[:-1])
        rew = 0.0
        if on.size > 0:
            rews = tolerance(goal_current[:-1][on] - self.normalized_state[on],
                             bounds=(0, _KEY_CLOSE_ENOUGH_TO_PRESSED), margin=_KEY_CLOSE_ENOUGH_TO_PRESSED * 10)
            rew += 0.5 * np.mean(rews)
        rew += 0.5 * (1 - float(failure))
        terms["key_press_reward"] = rew
        terms["sustain_reward"] = float(tolerance(goal_current[-1] - float(self.sustain_activation),
                                                  bounds=(0, _KEY_CLOSE_ENOUGH_TO_PRESSED),
                                                  margin=_KEY_CLOSE_ENOUGH_TO_PRESSED * 10))
        energy = 0.0
        hands = {None: (slice(0, 22), slice(22, 44)), "right": (slice(0, 22),), "left": (slice(22, 44),)}[self.side]
        for sl in hands:
            energy -= self.task._energy_penalty_coef * np.sum(np.abs(force[sl]) * np.abs(vel[sl]))
        terms["energy_reward"] = energy
        key_pos = lambda k: self.physics.geom_xpos[k] + self.key_geom_pos_offset[k]
        if self.has_fingering:
            dist = [np.linalg.norm(key_pos(k) - p.site_xpos[f]) for k, f in rh_cur]
            dist += [np.linalg.norm(key_pos(k) - p.site_xpos[5 + f]) for k, f in lh_cur]
            if not dist:
                terms["fingering_reward"] = 0.0
            else:
                terms["fingering_reward"] = float(np.mean(tolerance(
                    np.hstack(dist), bounds=(0, _FINGER_CLOSE_ENOUGH_TO_KEY), margin=_FINGER_CLOSE_ENOUGH_TO_KEY * 10)))
        elif self.side is None:
            if on.size == 0:
                terms["ot_fingering_reward"] = 1.0
            else:
                tips = np.concatenate([p.site_xpos[5:10], p.site_xpos[0:5]])  # left hand first
                kp = np.stack([key_pos(k) for k in on])
                dist = np.linalg.norm(tips[:, None] - kp[None], axis=-1)
                r, c = linear_sum_assignment(dist)
                terms["ot_fingering_reward"] = float(np.mean(tolerance(
                    dist[r, c], bounds=(0, _FINGER_CLOSE_ENOUGH_TO_KEY), margin=_FINGER_CLOSE_ENOUGH_TO_KEY * 10)))
        if not self.task._disable_forearm_reward:
            hit = any((c["geom1"] in self.rh_forearm and c["geom2"] in self.lh_forearm) or
                      (c["geom2"] in self.rh_forearm and c["geom1"] in self.lh_forearm)
                      for c in p.contacts() if c["dist"] <= 0)
            terms["forearm_reward"] = 0.0 if hit else 0.5
        self.reward_terms = terms
        reward = 0.0
        for v in terms.values():
            reward += v
        done = should_terminate
        if self.task._wrong_press_termination and failure and not done:
            self.discount = 0.0
            done = True
        elif self.task._wrong_press_termination and failure:
            pass  # should_terminate wins: discount stays 1 (piano_with_shadow_hands.py:213-220)
        return self.observation(), float(reward), bool(done)
