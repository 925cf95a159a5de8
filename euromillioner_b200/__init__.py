"""B200-native random-forest engine behind Spark ML's RandomForestClassifier surface."""
