-- Fixture written for the parser tests: the keys simulator.cpp:65-97 reads, with the
-- values of the reference's simulator config, in this repo's own words.
function Vector2(x, y) return {x = x, y = y} end
function Vector3(x, y, z)
  return {x = x, y = y, z = z}
end
function DegToRad(d)
  return math.pi * d / 180
end

--[[ block comment:
     init_config_file = "never evaluated" ]]
delta_t = 0.025
publish_tfs = true; publish_foot_to_base = true;
publish_map_to_odom = true

car_width = 0.281
car_length = 0.535
car_height = 0.15;
rear_axle_offset = -0.162
laser_loc = Vector3(0.2, 0, 0.15)

laser_noise_stddev = 0.01;
laser_angle_min = DegToRad(-135.0);
laser_angle_max = DegToRad(135.0);
laser_angle_increment = DegToRad(0.25);
laser_min_range = 0.4;
laser_max_range = 100.0;

local RobotType = {
    ACKERMANN_DRIVE="ACKERMANN_DRIVE",
    OMNIDIRECTIONAL_DRIVE="OMNIDIRECTIONAL_DRIVE",
    DIFF_DRIVE="DIFF_DRIVE"
}
robot_types = { RobotType.ACKERMANN_DRIVE, RobotType["DIFF_DRIVE"] }
laser_topic = "scan"
laser_frame = 'base_laser'
map_name = "maps/" .. "test" .. ".vectormap.txt"
start_poses = { {33.016267, 18.534157, 0}, Vector3(1.5, -2, DegToRad(90)) }
human_config_list = { "config/humans/h0.lua" ,
  "config/humans/h1.lua" }
