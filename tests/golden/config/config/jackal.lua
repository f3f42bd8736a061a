function Vector3(x, y, z) return {x = x, y = y, z = z} end
function DegToRad(d) return math.pi * d / 180 end
invert_linear_vel_cmds = false
invert_angular_vel_cmds = false
linear_pos_accel_limit = 3.0
linear_neg_accel_limit = 3.0
angular_pos_accel_limit = 3.0
angular_neg_accel_limit = 3.0
max_angular = 3.0
max_linear_vel = 2.0
rear_axle_offset = 0.0
min_turn_radius = 0.0
laser_loc = Vector3(0.07, 0, 0.5)
car_width = 0.43
car_length = 0.50
laser_angle_min = DegToRad(-180.0);
laser_angle_max = DegToRad(180.0);
laser_angle_increment = DegToRad(0.1);
laser_min_range = 0.4;
laser_max_range = 150.0;
