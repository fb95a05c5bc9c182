"""keras"""
