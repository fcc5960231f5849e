"""toric_rl_decoder_b200: B200-native (sm_100a) toric-code environment engine.

Drop-in for the env hot path of Lindeby/toric-RL-decoder: gym_ToricCode reset/step, the batched EnvSet
interface, generatePerspective* and the select-action input layout.  All compute is hand-written CUDA in
libtoric_b200.so (C ABI: include/toric_b200.h); importing the package does not need a GPU, using it does.
"""
from .buildlib import LIB_PATH, build  # noqa: F401
from .engine import (ACTOR_AUTO_RESET, ACTOR_RANDOM_ACTIONS, SAMPLE_DEPOLARIZING, SAMPLE_FIXED_PLUS_P,  # noqa: F401
                     SAMPLE_FIXED_WEIGHT, ToricEngine, p_to_thresh)
from .envs import EnvSetB200, ToricCodeB200, make  # noqa: F401
from .perspectives import (Action, Perspective, Transition, action_type, generatePerspective,  # noqa: F401
                           generatePerspectiveBatch, generatePerspectiveOptimized, transition_type)
from .policy import (calculateEpsilon, computePriorities, computePrioritiesParallel,  # noqa: F401
                     generatePerspectiveParallel, generateTransition, generateTransitionParallel,
                     predictMaxOptimized, rotate_state, selectAction, selectActionBatch, selectActionParallel,
                     select_action_device, shift_state, transitions_device)
from .hostpipe import HostActorPipeline, LazyStates  # noqa: F401
