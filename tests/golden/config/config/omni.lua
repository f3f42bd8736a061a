function DegToRad(d)
  return math.pi * d / 180
end
co_base_radius = 0.2
co_w0 = DegToRad(45.0)
co_max_angle_vel = math.pi
co_max_angle_accel = math.pi
co_max_speed = 1.2
co_max_accel = 3.0
