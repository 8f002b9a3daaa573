"""linear_programming_b200 -- B200-native dense LP engine."""
