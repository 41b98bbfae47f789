"""B200-native DQN learner hot path behind the dqn_zoo agent / replay API."""
__version__ = "0.1.0"
