"""bindings/python/Math/src/Constants.cpp"""
StandardAccelerationOfGravitation = 9.80665
