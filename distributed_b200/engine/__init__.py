"""engine"""
